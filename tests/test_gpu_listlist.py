"""GPU parity of the list x list operations and the on-disk formats: the reference's Boost tests through
the API (GV/test/testing_voxellist.cu:44-107, 341-460, 462-572), then randomised lists against the numpy oracle."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

DIMS = (89, 123, 74)
SV_START = 4


@pytest.fixture(scope="module")
def gv():
    import gpu_voxels_b200 as g
    g._lib.init()
    return g


def _box(oracle, mn, mx, delta):
    return oracle.create_box_of_points(mn, mx, delta)


def test_ref_collide_bitvoxellist_with_countingpermeaning_bitvoxellist(gv, oracle):
    """GV/test/testing_voxellist.cu:44-107: 50 swept-volume boxes against one obstacle box, per-meaning counts"""
    obstacle, swept = gv.BitVectorVoxelList(DIMS, 1.0), gv.BitVectorVoxelList(DIMS, 1.0)
    obstacle.insertMetaPointCloud(gv.MetaPointCloud([_box(oracle, (8, 8, 8), (11, 11, 11), 0.1)]), gv.eBVM_OCCUPIED)
    clouds, meanings = [], []
    for i in range(50):
        # "(max-min) * swept_ratio_delta * i + min" with int max/min and a float delta: float32 products, left to right
        # (30 * 0.02f * 15 = 8.9999995, not 9: that box reaches one voxel further down than exact arithmetic says)
        p = np.float32(np.float32(np.float32(30) * np.float32(0.02)) * np.float32(i)) + np.float32(0)
        clouds.append(_box(oracle, (p, p, p), (np.float32(p + np.float32(2.5)),) * 3, 0.1))
        meanings.append(SV_START + i)
    swept.insertMetaPointCloud(gv.MetaPointCloud(clouds), meanings)
    total, counts = swept.collideCountingPerMeaning(obstacle)
    got = " ".join("(%d|%d)" % (i, counts[i]) for i in range(256) if counts[i] > 0)
    assert got == "(14|1) (15|8) (16|8) (17|27) (18|27) (19|27) (20|8) (21|1) (22|1)"
    assert total == 108


def test_ref_bitvoxellist_findMatchingVoxels(gv, oracle):
    """GV/test/testing_voxellist.cu:341-410: lists of different length collided both ways"""
    side = 0.01
    l1, l2 = gv.BitVectorVoxelList(DIMS, side), gv.BitVectorVoxelList(DIMS, side)
    box1 = _box(oracle, (2.1, 0.1, 5.1), (4.9, 4.9, 9.9), 0.02)
    box2 = _box(oracle, (5.1, 5.1, 10.1), (9.9, 9.9, 14.9), 0.3)
    cp = [np.array([[np.float32(j / float(n) * 5.0), y, 1.0] for j in range(n)], np.float32) for n, y in ((40, 1.0), (30, 2.0), (20, 3.0))]
    l1.insertPointCloud(box1, 10)
    for c, mg in zip(cp, (11, 12, 13)):
        l1.insertPointCloud(c, mg)
    l2.insertPointCloud(box2, 20)
    for c, mg in zip(cp, (21, 22, 23)):
        l2.insertPointCloud(c, mg)
    assert len(l1) > len(l2)
    n, c = l1.collideCountingPerMeaning(l2)
    assert n == 90 and (c[11], c[12], c[13]) == (40, 30, 20) and c[21] == c[22] == c[23] == c[10] == c[20] == 0
    n, c = l2.collideCountingPerMeaning(l1)
    assert n == 90 and (c[21], c[22], c[23]) == (40, 30, 20) and c[11] == c[12] == c[13] == c[10] == c[20] == 0
    assert l1.collideWith(l2) == 90 and l2.collideWith(l1) == 90


def test_ref_bitvoxellist_bitshift_collision(gv):
    """GV/test/testing_voxellist.cu:412-460"""
    m1, m2 = gv.BitVectorVoxelList(DIMS, 1.0), gv.BitVectorVoxelList(DIMS, 1.0)
    pts = np.array([[0.3, 0.3, 0.3]], np.float32)
    shift_start = 50
    type_int = SV_START + shift_start
    for shift in range(0, shift_start + SV_START - 5):
        m1.clearMap(); m2.clearMap()
        m2.insertPointCloud(pts, type_int)
        m1.insertPointCloud(pts, type_int - shift)
        m2.shiftLeftSweptVolumeIDs(shift)
        assert m1.collideWithBitcheck(m2, 1) == (1 if shift <= shift_start else 0), shift


def _two_boxes(gv, oracle):
    return gv.MetaPointCloud([_box(oracle, (1.1, 1.1, 1.1), (3.9, 3.9, 3.9), 0.1), _box(oracle, (2.1, 2.1, 2.1), (4.9, 4.9, 4.9), 0.1)])


def test_ref_voxellist_equals_function(gv, oracle):
    """GV/test/testing_voxellist.cu:462-500"""
    l1, l2 = gv.BitVectorVoxelList(DIMS, 1.0), gv.BitVectorVoxelList(DIMS, 1.0)
    boxes = _two_boxes(gv, oracle)
    l1.insertMetaPointCloud(boxes, [11, 12]); l2.insertMetaPointCloud(boxes, [11, 12])
    assert l1.equals(l2)
    out = np.array([[9.22, 9.22, 9.22]], np.float32)
    l1.insertPointCloud(out, 11); l2.insertPointCloud(out, 12)
    assert not l1.equals(l2)


def test_ref_voxellist_disk_io(gv, oracle, tmp_path):
    """GV/test/testing_voxellist.cu:503-539, plus: the file is byte-identical to the reference's layout"""
    lst = gv.BitVectorVoxelList(DIMS, 1.0)
    lst.insertMetaPointCloud(_two_boxes(gv, oracle), [11, 12])
    path = os.path.join(str(tmp_path), "temp_list.lst")
    assert lst.writeToDisk(path)
    l2 = gv.BitVectorVoxelList(DIMS, 1.0)
    assert l2.readFromDisk(path)
    assert lst.equals(l2) and len(l2) > 0
    r = oracle.VoxelList(DIMS, 1.0)
    r.ids, r.coords, r.voxels = lst.download()
    assert open(path, "rb").read() == oracle.list_file_bytes(r)
    assert not gv.BitVectorVoxelList((89, 123, 75), 1.0).readFromDisk(path)   # dimensions differ
    assert not gv.BitVectorVoxelList(DIMS, 0.5).readFromDisk(path)            # side differs: an error for lists
    assert not l2.readFromDisk(path + ".missing")


def test_ref_bitvoxellist_subtract(gv, oracle):
    """GV/test/testing_voxellist.cu:541-572: two 27-voxel boxes overlapping in 8 voxels -> 19 remain"""
    l1, l2 = gv.BitVectorVoxelList(DIMS, 1.0), gv.BitVectorVoxelList(DIMS, 1.0)
    l1.insertPointCloud(_box(oracle, (1.1, 1.1, 1.1), (3.9, 3.9, 3.9), 0.1), 11)
    l2.insertPointCloud(_box(oracle, (2.1, 2.1, 2.1), (4.9, 4.9, 4.9), 0.1), 12)
    assert len(l1) == 27 and len(l2) == 27
    l1.subtract(l2)
    assert len(l1) == 19


def _random_list(gv, oracle, seed, dims, side, n_pts, n_meanings=5):
    rng = np.random.Generator(np.random.PCG64(seed))
    g, r = gv.BitVectorVoxelList(dims, side), oracle.VoxelList(dims, side)
    ext = np.array(dims) * side
    for k in range(n_meanings):
        pts = rng.uniform(0, ext * 0.7, size=(n_pts, 3)).astype(np.float32)
        mg = int(rng.integers(0, 256))
        g.insertPointCloud(pts, mg); r.insert_points(pts, mg)
    return g, r


@pytest.mark.parametrize("na,nb", [(30000, 4000), (4000, 30000), (1, 20000), (20000, 20000)])
def test_random_lists_match_oracle(gv, oracle, na, nb):
    dims, side = (64, 64, 40), 0.01
    a, ra = _random_list(gv, oracle, 200 + na % 7, dims, side, na)
    b, rb = _random_list(gv, oracle, 300 + nb % 7, dims, side, nb)
    assert np.array_equal(a.download()[0], ra.ids) and np.array_equal(b.download()[0], rb.ids)
    want = oracle.list_collide_list(ra, rb)
    assert a.collideWith(b) == want and (want > 0 or na == 1)
    tv = gv.BitVectorVoxel(); tv.bytes[:] = 0xFF  # must be overwritten, not OR-ed
    cnt, wt = oracle.list_collide_list_types(ra, rb)
    assert a.collideWithTypes(b, tv) == cnt and np.array_equal(tv.bytes, wt)
    total, counts = oracle.list_collide_list_per_meaning(ra, rb)
    gt, gc = a.collideCountingPerMeaning(b)
    assert gt == total and np.array_equal(gc, counts)
    for margin in (0, 1, 7, 31):
        assert a.collideWithBitcheck(b, margin) == oracle.list_collide_list_bitcheck(ra, rb, margin), margin
    assert a.collideWith(b, offset=(1, 0, 0)) == 0  # refused like in the reference
    # shift, then subtract
    a.shiftLeftSweptVolumeIDs(9); oracle.list_shift(ra, 9)
    assert np.array_equal(a.download()[2], ra.voxels)
    a.subtract(b); oracle.list_subtract(ra, rb)
    ids, co, vx = a.download()
    assert np.array_equal(ids, ra.ids) and np.array_equal(co, ra.coords) and np.array_equal(vx, ra.voxels)
    assert a.collideWith(b) == 0
    a.subtract(a)
    assert len(a) == 0


def _same(g, r):
    ids, co, vx = g.download()
    return np.array_equal(ids, r.ids) and np.array_equal(co, r.coords) and np.array_equal(vx, r.voxels)


@pytest.mark.parametrize("na,nb", [(30000, 4000), (4000, 30000), (0, 5000), (5000, 0)])
@pytest.mark.parametrize("off,new_meaning", [(None, None), ((3, -2, 5), None), ((-4, 1, 0), 77), ((0, 0, 0), 200),
                                             ((-70, 0, 0), None)])
def test_list_merge_matches_oracle(gv, oracle, na, nb, off, new_meaning):
    """TemplateVoxelList<BitVoxel>::merge (GV/voxellist/TemplateVoxelList.hpp:548-608): ids, coordinates (kept from the
    list's own entry for shared ids) and OR-merged voxels, with voxel offsets (incl. one that wraps ids below zero) and
    a replacing meaning"""
    dims, side = (64, 64, 40), 0.01
    a, ra = _random_list(gv, oracle, 210 + na % 7, dims, side, na) if na else (gv.BitVectorVoxelList(dims, side), oracle.VoxelList(dims, side))
    b, rb = _random_list(gv, oracle, 310 + nb % 7, dims, side, nb) if nb else (gv.BitVectorVoxelList(dims, side), oracle.VoxelList(dims, side))
    assert a.merge(b, off, new_meaning)
    oracle.list_merge(ra, rb, off if off is not None else (0, 0, 0), new_meaning)
    assert _same(a, ra) and len(a) == len(ra.ids)
    assert _same(b, rb)  # the merged list is untouched
    # a second merge of the same list changes nothing but the (already present) bits
    assert a.merge(b, off, new_meaning)
    oracle.list_merge(ra, rb, off if off is not None else (0, 0, 0), new_meaning)
    assert _same(a, ra)


def test_list_merge_reference_boxes_and_metric_offset(gv, oracle):
    """the reference's two 27-voxel boxes as two lists: 46 entries after the merge, 8 with both meanings; a metric offset
    is floor(c / side) per axis like mapToVoxelsSigned; a map cannot be merged into a list"""
    l1, l2 = gv.BitVectorVoxelList(DIMS, 1.0), gv.BitVectorVoxelList(DIMS, 1.0)
    l1.insertPointCloud(_box(oracle, (1.1, 1.1, 1.1), (3.9, 3.9, 3.9), 0.1), 11)
    l2.insertPointCloud(_box(oracle, (2.1, 2.1, 2.1), (4.9, 4.9, 4.9), 0.1), 12)
    assert l1.merge(l2) and len(l1) == 46
    bits = np.unpackbits(l1.download()[2], axis=1, bitorder="little")
    assert int((bits[:, 11] & bits[:, 12]).sum()) == 8
    l3 = gv.BitVectorVoxelList(DIMS, 1.0)
    assert l3.merge(l2, metric_offset=(10.5, -0.5, 2.0))  # -> voxel offset (10, -1, 2)
    r2, r3 = oracle.VoxelList(DIMS, 1.0), oracle.VoxelList(DIMS, 1.0)
    r2.insert_points(_box(oracle, (2.1, 2.1, 2.1), (4.9, 4.9, 4.9), 0.1), 12)
    oracle.list_merge(r3, r2, (10, -1, 2))
    assert _same(l3, r3) and l3.download()[1].min(axis=0).tolist() == [12, 1, 4]
    assert not l3.merge(gv.BitVectorVoxelMap(DIMS, 1.0))
    assert l3.merge(l3) and _same(l3, r3)  # into itself without offset: nothing changes


def test_list_list_edge_cases(gv, oracle):
    dims = (16, 16, 16)
    a, b = gv.BitVectorVoxelList(dims, 0.1), gv.BitVectorVoxelList(dims, 0.1)
    assert a.collideWith(b) == 0 and a.equals(b) and a.collideWithBitcheck(b, 3) == 0
    a.insertPointCloud(np.array([[0.05, 0.05, 0.05]], np.float32), 7)
    assert a.collideWith(b) == 0 and not a.equals(b)
    a.subtract(b)
    assert len(a) == 1
    c = gv.BitVectorVoxelList((16, 16, 17), 0.1)
    assert a.collideWith(c) == gv.SSIZE_MAX if hasattr(gv, "SSIZE_MAX") else True
    with pytest.raises(gv.GvxError):
        a.collideWithBitcheck(a, 32)
    a.shiftLeftSweptVolumeIDs(57)  # logged, no change
    assert a.download()[2][0, 0] == 0x80


@pytest.mark.parametrize("t", ["bit", "prob"])
def test_map_disk_io_roundtrip_and_format(gv, oracle, tmp_path, t):
    """TemplateVoxelMap::writeToDisk / readFromDisk, GV/voxelmap/TemplateVoxelMap.hpp:666-799"""
    dims, side = (33, 20, 17), 0.02
    cls = gv.BitVectorVoxelMap if t == "bit" else gv.ProbVoxelMap
    m = cls(dims, side)
    rng = np.random.Generator(np.random.PCG64(5))
    pts = rng.uniform(0, 0.6, size=(5000, 3)).astype(np.float32)
    m.insertPointCloud(pts, 1)
    path = os.path.join(str(tmp_path), "map.bin")
    assert m.writeToDisk(path)
    data = m.download()
    assert open(path, "rb").read() == oracle.map_file_bytes(data.reshape(-1), 0 if t == "bit" else 4, dims, side)
    m2 = cls(dims, side)
    assert m2.readFromDisk(path)
    assert np.array_equal(m2.download(), data)
    other = gv.BitVectorVoxelMap(dims, side)
    other.insertPointCloud(pts, 1)
    assert m2.collideWith(other, 1.0) == m.collideWith(other, 1.0) > 0   # the summary was rebuilt by the read
    assert not (gv.ProbVoxelMap if t == "bit" else gv.BitVectorVoxelMap)(dims, side).readFromDisk(path)  # type differs
    assert not cls((33, 20, 18), side).readFromDisk(path)
    assert cls(dims, 0.5).readFromDisk(path)   # a different side length is only a warning in the reference
    assert not m.writeToDisk(os.path.join(str(tmp_path), "no_such_dir", "x.bin"))
