"""gvx_batch_robot_collide_noalloc (gpu_voxels_b200/csrc/gvx_noalloc.cu): the planner batch with ALL scratch in a static
__device__ array inside the library -- no cudaMalloc anywhere in the call; no
counterpart in the reference. Checked against the existing entry point gvx_batch_robot_collide (counts are integers:
compared with ==, the tolerance of every test here) on two calls in a row that share the scratch, and on the limits it
must refuse. Collected LAST: the file name sorts after every other test file. Passes
alone in a fresh process and after the other tests (it makes its own maps and restores every option it touches)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

BIT, PROB = 0, 4


@pytest.fixture(scope="module")
def gv():
    import gpu_voxels_b200 as g
    g._lib.init()
    return g


def _links(seed, n_links, n_per_link):
    rng = np.random.Generator(np.random.PCG64(seed))
    return [(rng.uniform(-0.12, 0.12, size=(n_per_link + 37 * k, 3))).astype(np.float32) for k in range(n_links)]


def _transforms(seed, n_cfg, n_links, extent):
    rng = np.random.Generator(np.random.PCG64(seed))
    T = np.zeros((n_cfg, n_links, 4, 4), np.float32)
    for c in range(n_cfg):
        for k in range(n_links):
            a = rng.uniform(-np.pi, np.pi)
            T[c, k] = np.eye(4, dtype=np.float32)
            T[c, k, :2, :2] = [[np.cos(a), -np.sin(a)], [np.sin(a), np.cos(a)]]
            T[c, k, :3, 3] = rng.uniform(0.1, extent - 0.1, 3)
    return T


def _noalloc(gv, env, links, meanings, T, mask, thr):
    """-> counts (n_cfg,) uint64 through gvx_batch_robot_collide_noalloc: device inputs, device output, no sync inside"""
    import torch
    L = gv._lib.load()
    xyz_d = torch.from_numpy(np.concatenate(links)).cuda()
    tr_d = torch.from_numpy(np.ascontiguousarray(T.reshape(len(T), len(links), 16))).cuda()
    counts_d = torch.full((len(T),), -1, dtype=torch.int64, device="cuda")   # garbage: the call must overwrite it
    torch.cuda.synchronize()
    sizes = np.asarray([len(l) for l in links], np.uint32)
    mg = np.asarray(meanings, np.int32)
    mk = None if mask is None else np.ascontiguousarray(mask.bytes, np.uint8)
    rc = L.gvx_batch_robot_collide_noalloc(env.handle, C.c_void_p(xyz_d.data_ptr()), sizes.ctypes.data_as(C.c_void_p),
                                           len(sizes), mg.ctypes.data_as(C.c_void_p), C.c_void_p(tr_d.data_ptr()), len(T),
                                           mk.ctypes.data_as(C.c_void_p) if mk is not None else None, thr,
                                           C.c_void_p(counts_d.data_ptr()))
    gv._lib.check(rc)
    gv._lib.check(L.gvx_sync())
    return counts_d.cpu().numpy().astype(np.uint64)


@pytest.mark.parametrize("env_type", [BIT, PROB])
def test_noalloc_equals_the_existing_entry_point_on_two_calls_in_a_row(gv, env_type):
    dims, side = (96, 80, 64), 0.01
    rng = np.random.Generator(np.random.PCG64(91))
    env = (gv.BitVectorVoxelMap if env_type == BIT else gv.ProbVoxelMap)(dims, side)
    g = np.arange(18)
    local = np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(-1, 3)
    for _ in range(12):
        o = rng.integers(0, [96 - 18, 80 - 18, 64 - 18])
        env.insertPointCloud((((local + o) + 0.5) * side).astype(np.float32), 1)
    links = _links(92, 5, 3000)
    meanings = [4, 5, 0, 7, 130]          # link 2 carries eBVM_FREE: collides only through a mask that holds bit 0
    mask = gv.BitVectorVoxel()
    for b in (0, 5, 130):
        mask.setBit(b)
    mpc = gv.MetaPointCloud(links)
    first = None
    # two calls in a row share the static scratch: the second has fewer, different configurations and another mask, so
    # stale "seen" bits or a stale environment bitmap of the first would show
    for T, mk, thr in ((_transforms(93, 21, 5, 0.64), None, 1.0), (_transforms(94, 9, 5, 0.64), mask, 0.5)):
        want = gv.batch_robot_collide(env, mpc, meanings, T, mk, thr)
        got = _noalloc(gv, env, links, meanings, T, mk, thr)
        assert np.array_equal(got, want), (got, want)
        assert want.max() > 0 and len(set(want.tolist())) > 3
        first = want if first is None else first
    # and once more with the first inputs: same answer as the first time
    assert np.array_equal(_noalloc(gv, env, links, meanings, _transforms(93, 21, 5, 0.64), None, 1.0), first)
    # the voxel-streaming configuration builds the bitmap in the scratch instead of using the map's summary
    gv.set_option(gv.GVX_OPT_USE_SUMMARY, 0)
    try:
        assert np.array_equal(_noalloc(gv, env, links, meanings, _transforms(93, 21, 5, 0.64), None, 1.0), first)
    finally:
        gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)
    env.destroy()


def test_noalloc_refuses_what_does_not_fit(gv):
    L = gv._lib.load()
    assert L.gvx_noalloc_scratch_bytes() == 16 << 20
    dims, side = (128, 128, 128), 0.01     # bitmap 256 KiB: 1 + 63 fit the 16 MiB, 1 + 64 do not
    env = gv.BitVectorVoxelMap(dims, side)
    env.insertPointCloud(np.random.Generator(np.random.PCG64(5)).uniform(0.3, 0.9, size=(20000, 3)).astype(np.float32), 1)
    links = _links(95, 2, 2000)
    ok = _noalloc(gv, env, links, [4, 5], _transforms(96, 63, 2, 1.28), None, 1.0)
    assert np.array_equal(ok, gv.batch_robot_collide(env, gv.MetaPointCloud(links), [4, 5], _transforms(96, 63, 2, 1.28)))
    with pytest.raises(gv.GvxError, match="bytes of scratch"):
        _noalloc(gv, env, links, [4, 5], _transforms(96, 64, 2, 1.28), None, 1.0)
    many = _links(97, 17, 10)
    with pytest.raises(gv.GvxError, match="maximum of 16"):
        _noalloc(gv, env, many, list(range(4, 21)), _transforms(98, 2, 17, 1.28), None, 1.0)
    env.destroy()
