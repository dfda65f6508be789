"""GPU: insertPointCloudFromFile (GV/GpuVoxelsMap.cpp:36-49, GV/GpuVoxels.cpp:427-440) through the C ABI -- a file from
tests/golden/files/ goes through the host loader and the sm_100a insert kernel; the map must equal (a) the CPU oracle's
insert of the points the REFERENCE's readers returned for that file (tests/golden/pointcloud_files.npz) and (b) a plain
insertPointCloud of the same points. Bit-exact."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FILES = os.path.join(ROOT, "tests", "golden", "files")
GOLD = os.path.join(ROOT, "tests", "golden", "pointcloud_files.npz")
BIT, PROB = 0, 4
DIMS = (64, 48, 40)


@pytest.fixture(scope="module")
def gv():
    import gpu_voxels_b200 as g
    g._lib.init()
    return g


def _golden(name, shift, scaling):
    g = np.load(GOLD)
    for k in range(int(g["n_cases"])):
        a = g["args_%d" % k]
        if str(g["seq_%d" % k]) == name and bool(a[0]) == shift and a[1] == scaling:
            return g["xyz_%d" % k], tuple(float(v) for v in a[2:5])
    raise KeyError(name)


@pytest.mark.parametrize("name,shift,scaling,side", [("cube.binvox", True, 0.01, 0.01), ("grid32.binvox", True, 0.04, 0.01),
                                                      ("simple.xyz", True, 0.5, 0.1), ("messy.xyz", False, 1.0, 1.0)])
@pytest.mark.parametrize("map_type", [BIT, PROB])
def test_insert_from_file_equals_reference_points_through_the_oracle(gv, oracle, monkeypatch, name, shift, scaling, side, map_type):
    monkeypatch.chdir(FILES)  # the format is detected by substring of the whole path: keep it the bare name
    pts, off = _golden(name, shift, scaling)
    meaning = gv.eBVM_OCCUPIED if map_type == PROB else 37
    m = (gv.ProbVoxelMap if map_type == PROB else gv.BitVectorVoxelMap)(DIMS, side)
    assert m.insertPointCloudFromFile(name, False, meaning, shift, off, scaling)
    want = oracle.new_map(map_type, DIMS)
    oracle.insert_points(want, map_type, DIMS, side, pts, meaning)
    got = m.download().reshape(-1).view(np.uint8)
    assert np.array_equal(got, want.reshape(-1).view(np.uint8))
    assert got.any()
    m2 = (gv.ProbVoxelMap if map_type == PROB else gv.BitVectorVoxelMap)(DIMS, side)
    m2.insertPointCloud(pts, meaning)
    assert np.array_equal(m2.download().reshape(-1).view(np.uint8), got)


def test_facade_and_list_and_failures(gv, monkeypatch):
    monkeypatch.chdir(FILES)
    # 0.77 / 32 per binvox cell, scaled to one 1 cm voxel per cell, centred in the voxel
    off, scaling = (0.005, 0.005, 0.005), 0.415584
    ok, pts = gv.PointcloudFileHandler.Instance().loadPointCloud("grid32.binvox", False, None, True, off, scaling)
    assert ok and len(pts) == 1697
    gvl = gv.GpuVoxels()
    gvl.initialize(64, 64, 64, 0.01)
    gvl.addMap(gv.MT_BITVECTOR_VOXELMAP, "m")
    assert gvl.insertPointCloudFromFile("m", "grid32.binvox", False, 12, True, off, scaling)
    assert not gvl.insertPointCloudFromFile("nope", "grid32.binvox", False, 12)
    assert not gvl.insertPointCloudFromFile("m", "bad_magic.binvox", False, 12)
    assert not gvl.insertPointCloudFromFile("m", "cloud.txt", False, 12)
    assert not gvl.insertPointCloudFromFile("m", "missing.xyz", False, 12)
    lst = gv.BitVectorVoxelList((64, 64, 64), 0.01)
    assert lst.insertPointCloudFromFile("grid32.binvox", False, 12, True, off, scaling)
    ref = gv.BitVectorVoxelList((64, 64, 64), 0.01)
    ref.insertPointCloud(pts, 12)
    a, b = lst.download(), ref.download()
    assert len(a[0]) == 1697 and all(np.array_equal(x, y) for x, y in zip(a, b))
    # every list voxel hits the map built from the same file
    assert lst.collideWith(gvl.getMap("m")) == len(a[0])
    assert not lst.insertPointCloudFromFile("no_dim.binvox", False, 12)
    # empty cloud: true, nothing inserted
    e = gv.BitVectorVoxelMap(DIMS, 0.01)
    assert e.insertPointCloudFromFile("empty.xyz", False, 5)
    assert not e.download().any()
