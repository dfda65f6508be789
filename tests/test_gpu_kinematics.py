"""GPU parity of the batched DH forward kinematics against the CPU oracle, which is pinned against the
reference's own KinematicLink.cu / Matrix4f product compiled for the host (tests/test_oracle_vs_ref_cpu.py and
tests/golden/dh_chain.npz). Bit-exact: the matrices feed a float->voxel floor, so one ulp can move a voxel."""
import numpy as np
import pytest

from test_gpu_batch import _env, _links

pytestmark = pytest.mark.gpu

BIT, PROB = 0, 4
HALF_PI = np.float32(np.pi / 2)


@pytest.fixture(scope="module")
def gv():
    import gpu_voxels_b200 as g
    g._lib.init()
    return g


def _chain(n_links, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    dh = np.zeros((n_links, 5))
    dh[:, 0] = rng.uniform(0.0, 0.2, n_links)                          # d
    dh[:, 1] = rng.choice([0.0, HALF_PI, -HALF_PI, 0.3], n_links)      # theta
    dh[:, 2] = rng.uniform(0.0, 0.25, n_links)                         # a
    dh[:, 3] = rng.choice([0.0, HALF_PI, -HALF_PI, np.float32(np.pi), 1.1], n_links)  # alpha
    dh[:, 4] = rng.choice([0, 0, 0, 1], n_links)                       # joint type
    return dh.astype(np.float32).astype(np.float64)


def _joint_values(n_cfg, n_links, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    jv = rng.uniform(-2 * np.pi, 2 * np.pi, size=(n_cfg, n_links)).astype(np.float32)
    sp = np.array([0.0, -0.0, HALF_PI, -HALF_PI, np.pi, 1e-30, -1e-30, 1e5, -12345.678, 3e38, 1e-45, 0.78539816], np.float32)
    k = min(len(sp), n_cfg)
    for l in range(n_links):
        jv[:k, l] = np.roll(sp, l)[:k]
    return jv


def _bits(a):
    """bit patterns, with every NaN mapped to one pattern: x86 produces the default NaN 0xFFC00000 for inf * 0, the
    device the canonical 0x7FFFFFFF; a NaN matrix entry sends the point to voxel 0 on both (saturating cast)"""
    b = np.ascontiguousarray(a, np.float32).view(np.uint32).copy()
    b[np.isnan(np.ascontiguousarray(a, np.float32))] = 0x7FC00000
    return b


@pytest.mark.parametrize("n_links,n_cfg", [(1, 1), (7, 1024), (6, 100003), (33, 257)])
def test_dh_chain_matches_oracle_bitwise(gv, oracle, n_links, n_cfg):
    dh = _chain(n_links, 40 + n_links)
    jv = _joint_values(n_cfg, n_links, 41)
    got, amb = gv.dh_chain_transforms(dh, jv, return_ambiguous=True)
    want = oracle.dh_chain(dh[:, :4], dh[:, 4].astype(np.int32), jv)
    assert got.shape == want.shape == (n_cfg, n_links, 4, 4)
    bad = np.flatnonzero((_bits(got) != _bits(want)).reshape(n_cfg, -1).any(axis=1))
    assert bad.size == 0, "configurations %s differ (ambiguous reported: %d)" % (bad[:8], amb)
    assert amb <= max(4, n_cfg * n_links // 100000), "the bracket flags far more values than its width predicts"
    assert np.array_equal(got[:, 0], np.broadcast_to(np.eye(4, dtype=np.float32), (n_cfg, 4, 4)))  # link 0: identity


@pytest.mark.parametrize("bits", [30, 24, 8])
def test_host_recomputation_of_unproven_roundings_is_identical(gv, oracle, bits):
    """A wide bracket sends many (bits=30), then all (24, 8) configurations through the host path, beyond the size of
    the ambiguous list (4096): the result must not change."""
    dh = _chain(7, 50)
    jv = _joint_values(20000, 7, 51)
    want = oracle.dh_chain(dh[:, :4], dh[:, 4].astype(np.int32), jv)
    gv.set_option(gv.GVX_OPT_DH_CHECK_BITS, bits)
    try:
        got, amb = gv.dh_chain_transforms(dh, jv, return_ambiguous=True)
    finally:
        gv.set_option(gv.GVX_OPT_DH_CHECK_BITS, 50)
    assert amb > (100 if bits == 30 else 15000)
    assert np.array_equal(_bits(got), _bits(want))


def test_dh_errors(gv):
    dh = _chain(3, 1)
    bad = dh.copy(); bad[1, 4] = 2
    with pytest.raises(gv.GvxError):
        gv.dh_chain_transforms(bad, np.zeros((2, 3), np.float32))
    with pytest.raises(gv.GvxError):
        gv.set_option(gv.GVX_OPT_DH_CHECK_BITS, 60)
    assert gv.dh_chain_transforms(dh, np.zeros((0, 3), np.float32)).shape == (0, 3, 4, 4)


@pytest.mark.parametrize("t", [BIT, PROB])
def test_batch_collide_with_device_fk_equals_host_transforms(gv, oracle, t):
    dims, side = (96, 96, 96), 0.01
    env, r_env = _env(gv, oracle, t, dims, side, 61)
    fill = np.random.Generator(np.random.PCG64(65)).uniform(0, 0.96, size=(250000, 3)).astype(np.float32)
    env.insertPointCloud(fill, gv.eBVM_OCCUPIED)
    oracle.insert_points(r_env, t, dims, side, fill, 1)
    n_links = 6
    links = _links(62, n_links, 3000)
    dh = _chain(n_links, 63)
    dh[:, 0] += 0.05
    jv = _joint_values(96, n_links, 64)
    jv[:, 0] = np.abs(jv[:, 0]) % 1.5
    meanings = list(range(10, 10 + n_links))
    T = oracle.dh_chain(dh[:, :4], dh[:, 4].astype(np.int32), jv)
    sizes = [len(c) for c in links]
    want = oracle.batch_robot_collide(r_env, t, dims, side, np.concatenate(links), sizes, meanings, T)
    mpc = gv.MetaPointCloud(links)
    got_host = gv.batch_robot_collide(env, mpc, meanings, T)
    got_dev = gv.batch_robot_collide_dh(env, mpc, meanings, dh, jv)
    assert np.array_equal(got_host, want)
    assert np.array_equal(got_dev, want)
    assert want.sum() > 0


def test_links_without_cloud_and_clouds_without_link(gv, oracle):
    """KinematicChain.cu:104-124: a link without a cloud still multiplies into the chain; a cloud whose name matches no
    link keeps its coordinates."""
    dims, side = (64, 64, 64), 0.01
    env, r_env = _env(gv, oracle, BIT, dims, side, 71)
    links = _links(72, 4, 2000)          # clouds 0..3; cloud 3 has no link (static attachment)
    links[3] = (links[3] + np.float32(0.3)).astype(np.float32)
    dh = _chain(5, 73)                   # links 0..4; links 1 and 3 carry no cloud
    link_cloud = [0, -1, 1, -1, 2]
    jv = _joint_values(40, 5, 74)
    T5 = oracle.dh_chain(dh[:, :4], dh[:, 4].astype(np.int32), jv)
    T = np.zeros((40, 4, 4, 4), np.float32)
    T[:, 0], T[:, 1], T[:, 2] = T5[:, 0], T5[:, 2], T5[:, 4]
    T[:, 3] = np.eye(4, dtype=np.float32)
    meanings = [4, 5, 6, 7]
    want = oracle.batch_robot_collide(r_env, BIT, dims, side, np.concatenate(links), [len(c) for c in links], meanings, T)
    got = gv.batch_robot_collide_dh(env, gv.MetaPointCloud(links), meanings, dh, jv, link_cloud=link_cloud)
    assert np.array_equal(got, want)
    assert want.sum() > 0
