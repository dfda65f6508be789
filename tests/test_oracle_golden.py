"""The CPU oracle against every known-answer test the reference holds for the hot path.

Each case restates one Boost test of the reference (file:line in the docstring) with the reference's
default fixture (dims 89x123x74, 10000 points; GV/test/testing_fixtures.hpp:131-166) and checks the
oracle reproduces the expected integers. No GPU needed.
"""
import numpy as np
import pytest

DIMS = (89, 123, 74)
N = DIMS[0] * DIMS[1] * DIMS[2]
NPTS = 10000
BIT, PROB = 0, 4


def test_offset(oracle):
    """GV/test/testing_voxelmap.cu:58-92 (voxelmap/offset)"""
    sdim = (15, 16, 17)
    for off, truth in [((0, -2, 0), -30), ((1, 1, 1), 256), ((1, -2, -3), -749), ((3, -2, -1), -267),
                       ((-1, -2, 3), 689)]:
        assert oracle.voxel_index_signed(sdim, off) == truth


def test_float_to_probability(oracle):
    """GV/voxel/DefaultCollider.hpp:94-98: thr 1.0 -> 127, 0.5 -> 0, 0.1 -> -101 (truncation toward 0)"""
    assert oracle.float_to_probability(1.0) == 127
    assert oracle.float_to_probability(0.5) == 0
    assert oracle.float_to_probability(0.1) == -101
    assert oracle.float_to_probability(0.0) == -127


def test_collision(oracle):
    """GV/test/testing_voxelmap.cu:95-114 (voxelmap/collision): Prob x Prob, thr 0.1 -> 10000"""
    pts = oracle.create_equidistant_points(NPTS, DIMS, 1.0)
    assert len(pts) == NPTS
    m1, m2 = oracle.new_map(PROB, DIMS), oracle.new_map(PROB, DIMS)
    assert not oracle.insert_points(m1, PROB, DIMS, 1.0, pts, 1)
    oracle.insert_points(m2, PROB, DIMS, 1.0, pts, 1)
    assert oracle.collide_map_map(m1, PROB, m2, PROB, N, 0.1) == NPTS


def test_collision_with_offset(oracle):
    """GV/test/testing_voxelmap.cu:118-141: 8, then 18 with (-1,0,-1), 18 reversed with (1,0,1)"""
    p1 = oracle.create_box_of_points((2.1, 2.1, 2.1), (4.1, 4.1, 4.1), 0.5)
    p2 = oracle.create_box_of_points((3.1, 3.1, 3.1), (5.1, 5.1, 5.1), 0.5)
    m1, m2 = oracle.new_map(PROB, DIMS), oracle.new_map(PROB, DIMS)
    oracle.insert_points(m1, PROB, DIMS, 1.0, p1, 1)
    oracle.insert_points(m2, PROB, DIMS, 1.0, p2, 1)
    assert oracle.collide_map_map(m1, PROB, m2, PROB, N, 0.1) == 8
    assert oracle.collide_map_map(m1, PROB, m2, PROB, N, 0.1, oracle.voxel_index_signed(DIMS, (-1, 0, -1))) == 18
    assert oracle.collide_map_map(m2, PROB, m1, PROB, N, 0.1, oracle.voxel_index_signed(DIMS, (1, 0, 1))) == 18


def test_no_collision(oracle):
    """GV/test/testing_voxelmap.cu:143-167: 3-D checkerboard -> 0"""
    black, white = oracle.create_equidistant_points(NPTS, DIMS, 1.0, checkerboard=True)
    m1, m2 = oracle.new_map(PROB, DIMS), oracle.new_map(PROB, DIMS)
    oracle.insert_points(m1, PROB, DIMS, 1.0, black, 1)
    oracle.insert_points(m2, PROB, DIMS, 1.0, white, 1)
    assert oracle.collide_map_map(m1, PROB, m2, PROB, N, 0.1) == 0


CUBE8 = np.array([[x, y, z] for x in (2, 3) for y in (2, 3) for z in (2, 3)], np.float32)


def test_collision_with_types_probab(oracle):
    """GV/test/testing_voxelmap.cu:312-360: count 8, mask == exactly bits {14, 46, 103}"""
    m1, m2 = oracle.new_map(BIT, DIMS), oracle.new_map(PROB, DIMS)
    for mg in (4 + 10, 4 + 42, 4 + 99):
        oracle.insert_points(m1, BIT, DIMS, 1.0, CUBE8, mg)
    oracle.insert_points(m2, PROB, DIMS, 1.0, CUBE8, 1)
    cnt, mask = oracle.collide_bit_prob_types(m1, m2, N, 1.0)
    assert cnt == 8
    assert oracle.bits_of(mask) == [14, 46, 103]
    # side effect (VoxelMapOperations.hpp:211): colliding voxels of map_1 now carry eBVM_COLLISION
    cnt2, mask2 = oracle.collide_bit_prob_types(m1, m2, N, 1.0)
    assert cnt2 == 8 and oracle.bits_of(mask2) == [2, 14, 46, 103]


def test_bit_layout(oracle):
    """GV/test/testing_voxelmap.cu:364-398 + testing_bitvector.cu:48-51,109-112: bit k lives in byte k>>3 at k&7"""
    m = oracle.new_map(BIT, (1, 1, 1))
    for mg in (0, 2, 3, 4, 111, 254, 255):
        oracle.insert_points(m, BIT, (1, 1, 1), 1.0, [[0.5, 0.5, 0.5]], mg)
    assert m[0] == 29 and m[13] == 128 and m[31] == 192
    assert oracle.bits_of(m) == [0, 2, 3, 4, 111, 254, 255]
    truth = "10111" + "0" * 106 + "1" + "0" * 142 + "11"
    assert "".join("1" if (m[k >> 3] >> (k & 7)) & 1 else "0" for k in range(256)) == truth


def _two_boxes(oracle):
    b1 = oracle.create_box_of_points((1.1, 1.1, 1.1), (3.9, 3.9, 3.9), 0.1)
    b2 = oracle.create_box_of_points((2.1, 2.1, 2.1), (4.9, 4.9, 4.9), 0.1)
    return np.concatenate([b1, b2]), [len(b1), len(b2)]


def test_list_insert_metapointcloud(oracle):
    """GV/test/testing_voxellist.cu:282-338: 46 unique voxels, 8 with both bits {11,12}, all in the centre"""
    xyz, sizes = _two_boxes(oracle)
    lst = oracle.VoxelList(DIMS, 1.0)
    lst.insert_meta(xyz, sizes, [11, 12])
    assert len(lst) == 46
    assert np.all(np.diff(lst.ids.astype(np.int64)) > 0)
    both = [i for i in range(46) if oracle.bits_of(lst.voxels[i]) == [11, 12]]
    assert len(both) == 8
    assert all(2 <= c <= 3 for i in both for c in lst.coords[i])


@pytest.mark.parametrize("map_type", [PROB, BIT])
def test_list_collide_with_map(oracle, map_type):
    """GV/test/testing_voxellist.cu:109-145 (46), :149-236 (types: 46, mask == {11,12}) for both map types"""
    xyz, sizes = _two_boxes(oracle)
    lst = oracle.VoxelList(DIMS, 1.0)
    lst.insert_meta(xyz, sizes, [11, 12])
    m = oracle.new_map(map_type, DIMS)
    oracle.insert_meta(m, map_type, DIMS, 1.0, xyz, sizes, [1])
    cnt, mask, _ = lst.collide(1, m, map_type, 1.0, mark=False)
    assert cnt == 46 and oracle.bits_of(mask) == [11, 12]
    cnt, _, _ = lst.collide(0, m, map_type, 1.0)
    assert cnt == 46
    # per-meaning counts over the same colliding set: 27 voxels carry 11, 27 carry 12 (8 carry both);
    # every entry also carries eBVM_COLLISION from the marking collide above
    tot, _, per = lst.collide(3, m, map_type, 1.0, mark=False)
    assert per[11] == 27 and per[12] == 27 and per[2] == 46 and tot == 100


def test_list_collide_shifting(oracle):
    """GV/test/testing_voxellist.cu:239-279: x-offset 0..3 -> 27, 18, 9, 0"""
    b1 = oracle.create_box_of_points((1.1, 1.1, 1.1), (3.9, 3.9, 3.9), 0.1)
    lst = oracle.VoxelList(DIMS, 1.0)
    lst.insert_meta(b1, [len(b1)], [1])
    m = oracle.new_map(PROB, DIMS)
    oracle.insert_meta(m, PROB, DIMS, 1.0, b1, [len(b1)], [1])
    shift = np.float32(0.0)
    while shift < 4.0:
        off = oracle.voxel_index_signed(DIMS, (int(shift), 0, 0))
        cnt, _, _ = lst.collide(0, m, PROB, 1.0, off)
        assert cnt == (27 if shift < 1 else 18 if shift < 2 else 9 if shift < 3 else 0)
        shift = np.float32(shift + np.float32(0.5))


def test_bitmasked_collision(oracle):
    """GV/test/testing_voxellist.cu:739-864 (voxellists/bitmasked_collision), side 0.025"""
    side = 0.025
    lst = oracle.VoxelList(DIMS, side)
    m = oracle.new_map(PROB, DIMS)
    oracle.insert_points(m, PROB, DIMS, side, oracle.create_box_of_points((0.111, 0.111, 0.111), (0.799, 0.399, 0.399), 0.005), 1)
    for mn, mx, mg in [((0.111, 0.311, 0.111), (0.399, 0.599, 0.399), 34),
                       ((0.311, 0.311, 0.111), (0.599, 0.599, 0.399), 63),
                       ((0.511, 0.311, 0.111), (0.799, 0.599, 0.399), 102)]:
        lst.insert_points(oracle.create_box_of_points(mn, mx, 0.005), mg)
    for off3, expect in [((0, 0, 0), (0, 0, 9, 15, 21)), ((0, -4, 0), (0, 0, 18, 30, 42)),
                         ((0, -4, -4), (0, 0, 12, 20, 28))]:
        off = oracle.voxel_index_signed(DIMS, off3)
        mask = np.zeros(32, np.uint8)
        got = [lst.collide(2, m, PROB, 1.0, off, mask_in=mask)[0]]
        for bit in (33, 34, 63, 102):
            mask[bit >> 3] |= 1 << (bit & 7)
            got.append(lst.collide(2, m, PROB, 1.0, off, mask_in=mask)[0])
        assert got == [e * 64 for e in expect]


SV_START = 4  # eBVM_SWEPT_VOLUME_START


def test_collision_bitshift(oracle):
    """GV/test/testing_voxelmap.cu:173-229 (voxelmap/collision_bitshift): 8 voxels of meaning SV_START+40 against the
    same voxels of meaning i, window 20: collisions (all 8, only bit type_1) iff |i - type_1| <= 20"""
    pts = np.array([[x, y, z] for x in (2, 3) for y in (2, 3) for z in (2, 3)], np.float32)
    type_1, window = SV_START + 40, 20
    m1 = oracle.new_map(BIT, DIMS)
    oracle.insert_points(m1, BIT, DIMS, 1.0, pts, type_1)
    types = np.zeros(32, np.uint8)  # the reference never resets types_voxel between iterations
    for i in range(type_1 - window - 9, type_1 + window + 10):
        m2 = oracle.new_map(BIT, DIMS)
        oracle.insert_points(m2, BIT, DIMS, 1.0, pts, i)
        cnt, types = oracle.collide_bitcheck(m1, m2, N, window, 0, True, types)
        if cnt > 0:
            types[type_1 >> 3] &= ~np.uint8(1 << (type_1 & 7))
            assert not types.any()
        assert cnt == (0 if abs(i - type_1) > window else len(pts)), i


def test_bitvoxelmap_bitshift(oracle):
    """GV/test/testing_voxelmap.cu:238-304 (voxelmap/bitvoxelmap_bitshift): map_2 holds SV_START+50, map_1 holds
    SV_START+50-shift; after shifting map_2 by `shift` all 4 voxels collide (window 1) iff shift <= 50"""
    pts = np.array([[0.3] * 3, [1.3] * 3, [2.3] * 3, [3.3] * 3], np.float32)
    shift_start = 50
    type_int = SV_START + shift_start
    for shift in range(0, shift_start + SV_START - 5):
        m1, m2 = oracle.new_map(BIT, DIMS), oracle.new_map(BIT, DIMS)
        oracle.insert_points(m2, BIT, DIMS, 1.0, pts, type_int)
        oracle.insert_points(m1, BIT, DIMS, 1.0, pts, type_int - shift)
        assert oracle.shift_map(m2, N, shift)
        cnt, _ = oracle.collide_bitcheck(m1, m2, N, 1, 0, True)
        assert cnt == (len(pts) if shift <= shift_start else 0), shift
    assert not oracle.shift_map(m2, N, 57)


def _olist(oracle, side=1.0):
    return oracle.VoxelList(DIMS, side)


def test_list_list_countingpermeaning(oracle):
    """GV/test/testing_voxellist.cu:44-107: "(14|1) (15|8) (16|8) (17|27) (18|27) (19|27) (20|8) (21|1) (22|1)", 108"""
    obstacle, swept = _olist(oracle), _olist(oracle)
    obstacle.insert_points(oracle.create_box_of_points((8, 8, 8), (11, 11, 11), 0.1), 1)
    clouds, meanings = [], []
    for i in range(50):
        p = np.float32(np.float32(np.float32(30) * np.float32(0.02)) * np.float32(i)) + np.float32(0)
        clouds.append(oracle.create_box_of_points((p, p, p), (np.float32(p + np.float32(2.5)),) * 3, 0.1))
        meanings.append(SV_START + i)
    swept.insert_meta(np.concatenate(clouds), [len(c) for c in clouds], meanings)
    total, counts = oracle.list_collide_list_per_meaning(swept, obstacle)
    got = " ".join("(%d|%d)" % (i, counts[i]) for i in range(256) if counts[i] > 0)
    assert got == "(14|1) (15|8) (16|8) (17|27) (18|27) (19|27) (20|8) (21|1) (22|1)" and total == 108


def test_list_list_find_matching_voxels(oracle):
    """GV/test/testing_voxellist.cu:341-410"""
    l1, l2 = _olist(oracle, 0.01), _olist(oracle, 0.01)
    cp = [np.array([[np.float32(j / float(n) * 5.0), y, 1.0] for j in range(n)], np.float32) for n, y in ((40, 1.0), (30, 2.0), (20, 3.0))]
    l1.insert_points(oracle.create_box_of_points((2.1, 0.1, 5.1), (4.9, 4.9, 9.9), 0.02), 10)
    l2.insert_points(oracle.create_box_of_points((5.1, 5.1, 10.1), (9.9, 9.9, 14.9), 0.3), 20)
    for c, m1, m2 in zip(cp, (11, 12, 13), (21, 22, 23)):
        l1.insert_points(c, m1); l2.insert_points(c, m2)
    n, c = oracle.list_collide_list_per_meaning(l1, l2)
    assert n == 90 and (c[11], c[12], c[13]) == (40, 30, 20) and c[21] == c[10] == 0
    n, c = oracle.list_collide_list_per_meaning(l2, l1)
    assert n == 90 and (c[21], c[22], c[23]) == (40, 30, 20) and c[11] == c[20] == 0


def test_list_list_bitshift_collision_and_subtract(oracle):
    """GV/test/testing_voxellist.cu:412-460 (window 1 after shifting) and :541-572 (27 + 27 boxes, 8 shared -> 19)"""
    pts = np.array([[0.3, 0.3, 0.3]], np.float32)
    for shift in range(0, 50 + SV_START - 5):
        m1, m2 = _olist(oracle), _olist(oracle)
        m2.insert_points(pts, SV_START + 50)
        m1.insert_points(pts, SV_START + 50 - shift)
        oracle.list_shift(m2, shift)
        assert oracle.list_collide_list_bitcheck(m1, m2, 1) == (1 if shift <= 50 else 0), shift
    l1, l2 = _olist(oracle), _olist(oracle)
    l1.insert_points(oracle.create_box_of_points((1.1, 1.1, 1.1), (3.9, 3.9, 3.9), 0.1), 11)
    l2.insert_points(oracle.create_box_of_points((2.1, 2.1, 2.1), (4.9, 4.9, 4.9), 0.1), 12)
    oracle.list_subtract(l1, l2)
    assert len(l1) == 19


def test_update_logodds_openmp_loop_equals_the_sequential_loop(oracle):
    """bench.py's CPU baseline folds frames with gvo_update_logodds_omp (all host threads, byte CAS): it has to be the
    sequential restatement byte for byte -- dense clumps (saturation at 127), NaN points, the z < 0 re-projection branch,
    negative deltas (saturation at -127), z-slabs, 1 and many threads."""
    dims, side = (64, 48, 40), 0.02
    rng = np.random.Generator(np.random.PCG64(9))
    pts = rng.uniform(-0.7, 0.7, size=(120000, 3)).astype(np.float32)
    pts[:50000] *= np.float32(0.03)
    pts[1000:1100] = np.nan
    pose = np.eye(4, dtype=np.float32)[:3].copy()
    pose[:, 3] = (0.6, 0.5, 0.4)
    before = oracle.num_threads()
    try:
        for threads in (1, 7, max(2, before)):
            oracle.set_num_threads(threads)
            for z0, z1 in ((0, 40), (11, 29)):
                a, b = oracle.new_map(4, dims, z0, z1), oracle.new_map(4, dims, z0, z1)
                for delta in (72, 72, -10, 72, -127, 30):
                    oracle.update_logodds(a, dims, side, pts, pose, delta, z0, z1)
                    oracle.update_logodds(b, dims, side, pts, pose, delta, z0, z1, parallel=True)
                    assert np.array_equal(a, b), (threads, z0, delta)
                assert (a == 127).sum() > 5 and (a == -128).sum() > 1000
    finally:
        oracle.set_num_threads(before)
