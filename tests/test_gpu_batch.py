"""GPU parity of the batched robot-configuration collision check (BASELINE.json configs[2]) against the CPU oracle,
which restates the planner's per-configuration clear / transform / insert / collideWith sequence literally."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

BIT, PROB = 0, 4


@pytest.fixture(scope="module")
def gv():
    import gpu_voxels_b200 as g
    g._lib.init()
    return g


def _links(seed, n_links, n_pts):
    rng = np.random.Generator(np.random.PCG64(seed))
    out = []
    for k in range(n_links):
        r, l = rng.uniform(0.02, 0.05), rng.uniform(0.08, 0.2)
        t, a = rng.uniform(0, l, n_pts), rng.uniform(0, 2 * np.pi, n_pts)
        out.append(np.stack([r * np.cos(a), r * np.sin(a), t], -1).astype(np.float32))
    return out


def _transforms(seed, n_cfg, n_links, extent):
    """random rigid transforms (fp32 rotation about a random axis + translation inside / slightly outside the map)"""
    rng = np.random.Generator(np.random.PCG64(seed))
    T = np.zeros((n_cfg, n_links, 4, 4), np.float32)
    for c in range(n_cfg):
        for k in range(n_links):
            ax = rng.normal(size=3); ax /= np.linalg.norm(ax)
            th = rng.uniform(-np.pi, np.pi)
            K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
            R = np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)
            T[c, k, :3, :3] = R.astype(np.float32)
            T[c, k, :3, 3] = rng.uniform(-0.05, extent + 0.05, size=3).astype(np.float32)
            T[c, k, 3, 3] = 1
    return T


def _env(gv, oracle, t, dims, side, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    pts = []
    ext = np.array(dims) * side
    for _ in range(12):
        lo = rng.uniform(0, ext * 0.8)
        hi = lo + rng.uniform(0.05, 0.2, size=3)
        pts.append(rng.uniform(lo, hi, size=(20000, 3)))
    pts = np.concatenate(pts).astype(np.float32)
    m = (gv.ProbVoxelMap if t == PROB else gv.BitVectorVoxelMap)(dims, side)
    r = oracle.new_map(t, dims)
    m.insertPointCloud(pts, gv.eBVM_OCCUPIED)
    oracle.insert_points(r, t, dims, side, pts, 1)
    return m, r


@pytest.mark.parametrize("t", [BIT, PROB])
@pytest.mark.parametrize("use_mask", [False, True])
def test_batch_matches_planner_sequence(gv, oracle, t, use_mask):
    dims, side = (64, 64, 64), 0.01
    env, r_env = _env(gv, oracle, t, dims, side, 61)
    links = _links(62, 5, 4000)
    meanings = [4, 5, 0, 7, 130]  # link 2 carries eBVM_FREE: never collides without a mask
    T = _transforms(63, 24, 5, 0.64)
    mask = gv.BitVectorVoxel()
    if use_mask:
        for b in (0, 5, 130):
            mask.setBit(b)
    got = gv.batch_robot_collide(env, gv.MetaPointCloud(links), meanings, T, mask if use_mask else None, 1.0)
    want = oracle.batch_robot_collide(r_env, t, dims, side, np.concatenate(links), [len(l) for l in links], meanings,
                                      T, mask.bytes if use_mask else None, 1.0)
    assert np.array_equal(got, want)
    assert want.max() > 0 and len(set(want.tolist())) > 3
    # env untouched
    assert np.array_equal(env.download().reshape(-1).view(np.uint8), r_env.view(np.uint8))


def test_batch_equals_per_configuration_api_calls(gv, oracle):
    """the batch result equals what the reference-style call sequence produces on the GPU path itself"""
    dims, side = (48, 40, 56), 0.02
    env, _ = _env(gv, oracle, BIT, dims, side, 71)
    links = _links(72, 3, 3000)
    meanings = [10, 11, 12]
    T = _transforms(73, 6, 3, 0.8)
    got = gv.batch_robot_collide(env, gv.MetaPointCloud(links), meanings, T)
    rob = gv.BitVectorVoxelMap(dims, side)
    for c in range(len(T)):
        rob.clearMap()
        rob.insertMetaPointCloud(gv.MetaPointCloud(links), meanings, transforms=T[c])
        assert rob.collideWith(env, 1.0) == got[c]


def test_batch_waves_with_tiny_scratch(gv, oracle):
    """scratch budget forces several waves of configurations; result unchanged"""
    dims, side = (64, 64, 64), 0.01
    env, r_env = _env(gv, oracle, BIT, dims, side, 81)
    links = _links(82, 2, 5000)
    T = _transforms(83, 37, 2, 0.64)
    L = gv._lib.load()
    full = gv.batch_robot_collide(env, gv.MetaPointCloud(links), [4, 9], T)
    gv._lib.check(L.gvx_batch_set_scratch_budget(5 * (64 ** 3 // 8)))  # 5 configurations per wave
    try:
        waved = gv.batch_robot_collide(env, gv.MetaPointCloud(links), [4, 9], T)
    finally:
        gv._lib.check(L.gvx_batch_set_scratch_budget(4 << 30))
    assert np.array_equal(full, waved)
    want = oracle.batch_robot_collide(r_env, BIT, dims, side, np.concatenate(links), [5000, 5000], [4, 9], T, None, 1.0)
    assert np.array_equal(full, want)
