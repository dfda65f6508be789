"""The numpy restatement of TemplateVoxelList<BitVoxel>::merge (oracle/gvoracle.py list_merge;
GV/voxellist/TemplateVoxelList.hpp:548-608) pinned against the part of the oracle that the reference's own test pins
(make_unique through insertMetaPointCloud, GV/test/testing_voxellist.cu:280-338): merging list(B) into list(A) must equal
inserting B's points into list(A); a voxel offset equals inserting the shifted points; a new meaning equals inserting
B's points with that meaning. No reference test exercises merge itself."""
import numpy as np

DIMS = (89, 123, 74)


def _lists_equal(a, b):
    return np.array_equal(a.ids, b.ids) and np.array_equal(a.coords, b.coords) and np.array_equal(a.voxels, b.voxels)


def _cloud(seed, n, lo, hi):
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.uniform(lo, hi, size=(n, 3)).astype(np.float32)


def test_reference_boxes_merged_as_lists(oracle):
    """the reference's 27 + 27 voxel boxes (8 shared): 46 entries, the shared ones carry both meanings"""
    b1 = oracle.create_box_of_points((1.1, 1.1, 1.1), (3.9, 3.9, 3.9), 0.1)
    b2 = oracle.create_box_of_points((2.1, 2.1, 2.1), (4.9, 4.9, 4.9), 0.1)
    a, b = oracle.VoxelList(DIMS, 1.0), oracle.VoxelList(DIMS, 1.0)
    a.insert_points(b1, 11)
    b.insert_points(b2, 12)
    oracle.list_merge(a, b)
    both = oracle.VoxelList(DIMS, 1.0)
    both.insert_meta(np.concatenate([b1, b2]), [len(b1), len(b2)], [11, 12])
    assert len(a) == 46 and _lists_equal(a, both)
    bits = np.unpackbits(a.voxels, axis=1, bitorder="little")
    assert int((bits[:, 11] & bits[:, 12]).sum()) == 8


def test_merge_equals_insert_with_offset_and_new_meaning(oracle):
    side = 0.5
    pa, pb = _cloud(1, 4000, 2.0, 20.0), _cloud(2, 3000, 5.0, 25.0)
    for off, new_meaning in (((0, 0, 0), None), ((3, -2, 5), None), ((-4, 1, 0), 77), ((0, 0, 0), 200)):
        a, b = oracle.VoxelList(DIMS, side), oracle.VoxelList(DIMS, side)
        a.insert_points(pa, 20)
        b.insert_meta(pb, [1000, 2000], [30, 31])
        oracle.list_merge(a, b, off, new_meaning)
        want = oracle.VoxelList(DIMS, side)
        want.insert_points(pa, 20)
        # the same voxels reached by points: centre of every voxel of b, moved by the offset
        centre = ((b.coords.astype(np.int64) + np.asarray(off)).astype(np.float32) + np.float32(0.5)) * np.float32(side)
        for bit in ([new_meaning] if new_meaning is not None else [30, 31]):
            sel = np.ones(len(b), bool) if new_meaning is not None else (b.voxels[:, bit >> 3] >> (bit & 7) & 1).astype(bool)
            want.insert_points(centre[sel], bit)
        assert _lists_equal(a, want), (off, new_meaning)
    # ids wrap like the reference's uint32 arithmetic when the offset leaves the grid
    a, b = oracle.VoxelList(DIMS, side), oracle.VoxelList(DIMS, side)
    b.insert_points(np.array([[0.1, 0.1, 0.1]], np.float32), 9)
    oracle.list_merge(a, b, (-1, 0, 0))
    assert a.ids.tolist() == [0xFFFFFFFF] and a.coords.tolist() == [[0xFFFFFFFF, 0, 0]]
