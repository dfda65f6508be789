"""GPU parity against the REFERENCE'S OWN CUDA KERNELS, compiled unmodified for sm_100a from /root/reference into
oracle/_ref/libgvref_cuda.so (oracle/ref_cuda_driver.cu, built in the authoring container; the .so travels to the GPU
box, the reference sources do not). This pins what no reference unit test pins: the device float->voxel rounding for
NaN / negative / huge / boundary coordinates, the FMA contraction of the transform, the sensor-frame re-projection, the
collision write-back and the list gather -- all compared bit for bit on the same inputs in the same process."""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "oracle", "_ref", "libgvref_cuda.so")
BIT, PROB = 0, 4


@pytest.fixture(scope="module")
def env():
    if not os.path.exists(SO):
        pytest.skip("oracle/_ref/libgvref_cuda.so not built (needs /root/reference at build time)")
    import torch
    import gpu_voxels_b200 as gv
    gv._lib.init()
    ref = C.CDLL(SO)
    ref.gvref_collide_map_map.restype = C.c_longlong
    ref.gvref_collide_list_map.restype = C.c_longlong
    return gv, ref, torch


def _p(t):
    return C.c_void_p(t.data_ptr())


def _specials(p):
    if len(p) < 16:
        return p
    p[5] = [np.nan, 0.5, 0.5]
    p[6] = [0.5, np.inf, 0.5]
    p[7] = [-np.inf, 0.5, 0.5]
    p[8] = [-0.3, 0.4, 0.4]
    p[9] = [1e20, 0.1, 0.1]
    p[10] = [-0.0, 0.0, 0.0]
    p[11] = [0.29999998, 0.3, 0.30000004]
    p[12] = [4294967296.0 * 0.01, 0.2, 0.2]  # exactly 2^32 voxels
    p[13] = [-1e-30, 1e-30, 0.635]
    return p


def _ref_map(torch, t, n):
    return torch.full((n,), -128, dtype=torch.int8, device="cuda") if t == PROB else \
        torch.zeros(n * 32, dtype=torch.uint8, device="cuda")


@pytest.mark.parametrize("t", [BIT, PROB])
@pytest.mark.parametrize("side", [0.01, 0.025, 0.0173])
def test_insert_points_float_to_index_matches_reference_kernel(env, t, side):
    gv, ref, torch = env
    dims = (64, 64, 64)
    n = 64 ** 3
    rng = np.random.Generator(np.random.PCG64(7))
    pts = _specials(rng.uniform(-0.05, 64 * side + 0.05, size=(200000, 3)).astype(np.float32))
    # points ON voxel boundaries k*side as fp32 products: the IEEE division must land exactly where the reference's does
    k = rng.integers(0, 64, size=(20000, 3)).astype(np.float32)
    pts = np.concatenate([pts, (k * np.float32(side)).astype(np.float32),
                          np.nextafter((k * np.float32(side)).astype(np.float32), np.float32(-1))])
    m = (gv.ProbVoxelMap if t == PROB else gv.BitVectorVoxelMap)(dims, side)
    meaning = 1 if t == PROB else 37
    out = m.insertPointCloud(pts, meaning)
    r = _ref_map(torch, t, n)
    pd = torch.from_numpy(pts).cuda()
    flag = C.c_int(0)
    assert ref.gvref_insert_points(_p(r), int(t == PROB), *dims, C.c_float(side), _p(pd), len(pts), meaning, C.byref(flag)) == 0
    torch.cuda.synchronize()
    assert bool(flag.value) == out
    got = m.download().reshape(-1).view(np.uint8)
    want = r.cpu().numpy().view(np.uint8)
    diff = np.flatnonzero(got != want)
    assert diff.size == 0, "boundary-voxel mismatches vs the reference kernel at bytes %s" % diff[:10]


def test_transform_cloud_matches_reference_kernel_bitwise(env):
    gv, ref, torch = env
    rng = np.random.Generator(np.random.PCG64(8))
    pts = rng.normal(size=(100000, 3)).astype(np.float32) * np.float32(3)
    M = np.eye(4, dtype=np.float32)
    a = rng.normal(size=3); a /= np.linalg.norm(a)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    M[:3, :3] = (np.eye(3) + np.sin(0.7) * K + (1 - np.cos(0.7)) * K @ K).astype(np.float32)
    M[:3, 3] = [0.123, -4.5, 7.25]
    got = gv.transform_points(pts, M)
    pd, od = torch.from_numpy(pts).cuda(), torch.empty(pts.shape, dtype=torch.float32, device="cuda")
    Mc = np.ascontiguousarray(M.reshape(16))
    assert ref.gvref_transform_cloud(Mc.ctypes.data_as(C.c_void_p), _p(pd), _p(od), len(pts)) == 0
    assert np.array_equal(got.view(np.uint32), od.cpu().numpy().view(np.uint32))


def test_insert_meta_matches_reference_kernel(env):
    gv, ref, torch = env
    dims, side = (96, 64, 80), 0.01
    n = dims[0] * dims[1] * dims[2]
    rng = np.random.Generator(np.random.PCG64(9))
    clouds = [_specials(rng.uniform(-0.02, 0.98, size=(sz, 3)).astype(np.float32)) for sz in (30000, 1, 50000, 777)]
    meanings = [4, 20, 41, 255]  # one byte each: the reference's non-atomic byte RMW cannot lose bits here
    m = gv.BitVectorVoxelMap(dims, side)
    out = m.insertMetaPointCloud(gv.MetaPointCloud(clouds), meanings)
    r = _ref_map(torch, BIT, n)
    pd = torch.from_numpy(np.concatenate(clouds)).cuda()
    sizes = np.array([len(c) for c in clouds], np.uint32)
    mg = np.array(meanings, np.int32)
    flag = C.c_int(0)
    assert ref.gvref_insert_meta(_p(r), 0, *dims, C.c_float(side), _p(pd), sizes.ctypes.data_as(C.c_void_p), len(sizes),
                                 mg.ctypes.data_as(C.c_void_p), C.byref(flag)) == 0
    assert bool(flag.value) == out
    assert np.array_equal(m.download().reshape(-1), r.cpu().numpy())


def test_sensor_update_matches_reference_kernels_where_they_are_race_free(env, oracle):
    """kernelTransformSensorData + kernelInsertSensorData: the reference's update is a non-atomic read-modify-write, so
    voxels hit by several pixels may lose increments there. Compared exactly on voxels hit once; on the others the
    reference may only be LOWER (lost updates), and the set of touched voxels must be identical."""
    gv, ref, torch = env
    dims, side = (128, 128, 64), 0.01
    n = dims[0] * dims[1] * dims[2]
    rng = np.random.Generator(np.random.PCG64(10))
    pts = rng.uniform(-0.6, 0.6, size=(60000, 3)).astype(np.float32)
    pts[rng.uniform(size=len(pts)) < 0.02] = np.nan
    pose = np.array([[0.8, -0.6, 0, 0.64], [0.6, 0.8, 0, 0.64], [0, 0, 1, 0.3]], np.float32)  # z<0 branch for pz < -0.3
    m = gv.ProbVoxelMap(dims, side)
    m.insertSensorData(pts, pose)
    r = _ref_map(torch, PROB, n)
    pd, tmp = torch.from_numpy(pts).cuda(), torch.empty(pts.shape, dtype=torch.float32, device="cuda")
    pc = np.ascontiguousarray(pose.reshape(12))
    assert ref.gvref_insert_sensor_data(_p(r), *dims, C.c_float(side), pc.ctypes.data_as(C.c_void_p), _p(pd), _p(tmp), len(pts)) == 0
    got, want = m.download(), r.cpu().numpy()
    hits = oracle.new_map(PROB, dims)
    oracle.update_logodds(hits, dims, side, pts, pose, 1)
    hits = hits.astype(np.int32) + 128
    assert (hits == 1).sum() > 10000 and (hits > 1).sum() > 100
    assert np.array_equal(got != -128, want != -128), "different set of touched voxels"
    assert np.array_equal(got[hits <= 1], want[hits <= 1]), "rounding difference on race-free voxels"
    assert np.all(want[hits > 1] <= got[hits > 1]) and np.all(want[hits > 1] >= -56)


@pytest.mark.parametrize("ta,tb", [(BIT, PROB), (PROB, BIT), (PROB, PROB), (BIT, BIT)])
def test_collide_map_map_matches_reference_kernel(env, ta, tb):
    gv, ref, torch = env
    dims, side = (128, 96, 64), 0.01
    n = dims[0] * dims[1] * dims[2]
    rng = np.random.Generator(np.random.PCG64(11))
    pa = rng.uniform(0, 0.96, size=(150000, 3)).astype(np.float32)
    pb = np.concatenate([pa[:40000], rng.uniform(0, 0.96, size=(110000, 3)).astype(np.float32)])
    a = (gv.ProbVoxelMap if ta == PROB else gv.BitVectorVoxelMap)(dims, side)
    b = (gv.ProbVoxelMap if tb == PROB else gv.BitVectorVoxelMap)(dims, side)
    pose = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0.001]], np.float32)
    for mp_, t, p in ((a, ta, pa), (b, tb, pb)):
        if t == PROB:
            mp_.insertSensorData(p, pose)
            mp_.insertSensorData(p[::2], pose)
        else:
            mp_.insertPointCloud(p, 9)
    for thr in (0.5, 0.3, 1.0):
        ra = torch.from_numpy(a.download().reshape(-1).view(np.uint8).copy()).cuda()
        rb = torch.from_numpy(b.download().reshape(-1).view(np.uint8).copy()).cuda()
        want = ref.gvref_collide_map_map(_p(ra), ta, _p(rb), tb, n, C.c_float(thr))
        got = a.collideWith(b, thr)
        assert got == want and want >= 0
        assert np.array_equal(a.download().reshape(-1).view(np.uint8), ra.cpu().numpy()), "collision write-back differs"
    assert got > 0 or ta == tb == PROB


@pytest.mark.parametrize("t", [PROB, BIT])
def test_collide_list_map_matches_reference_kernel(env, t):
    gv, ref, torch = env
    dims, side = (96, 80, 72), 0.01
    rng = np.random.Generator(np.random.PCG64(12))
    lst = gv.BitVectorVoxelList(dims, side)
    lst.insertMetaPointCloud(gv.MetaPointCloud([rng.uniform(0, 0.9, size=(60000, 3)).astype(np.float32) for _ in range(3)]),
                             [5, 40, 200])
    m = (gv.ProbVoxelMap if t == PROB else gv.BitVectorVoxelMap)(dims, side)
    m.insertPointCloud(rng.uniform(0, 0.9, size=(200000, 3)).astype(np.float32), 1)
    md = torch.from_numpy(m.download().reshape(-1).view(np.uint8).copy()).cuda()
    for off3 in ((0, 0, 0), (2, -1, 3), (0, 0, -71), (-95, 0, 0)):
        ids, _, vx = lst.download()
        idd, vxd = torch.from_numpy(ids).cuda(), torch.from_numpy(vx.copy()).cuda()
        want = ref.gvref_collide_list_map(_p(idd), _p(vxd), len(ids), _p(md), t, *dims, C.c_float(1.0), *off3)
        got = lst.collideWith(m, 1.0, off3)
        assert got == want and want >= 0
        assert np.array_equal(lst.download()[2], vxd.cpu().numpy()), "eBVM_COLLISION marks on the list differ"
