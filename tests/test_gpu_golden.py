"""The CUDA path against the golden vectors of tests/golden/ (outputs of the reference's own code, see
tests/golden/make_golden.py) -- directly, without the oracle in between."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
BIT, PROB = 0, 4


@pytest.fixture(scope="module")
def gv():
    import gpu_voxels_b200 as g
    g._lib.init()
    return g


def _load(name):
    return np.load(os.path.join(G, name), allow_pickle=False)


def _map(gv, t, dims, data):
    m = (gv.ProbVoxelMap if t == PROB else gv.BitVectorVoxelMap)(dims, 0.01)
    m.upload(data)
    return m


@pytest.mark.parametrize("use_summary", [1, 0])
def test_collide_map_map_golden(gv, use_summary):
    z = _load("colliders.npz")
    dims = tuple(int(v) for v in z["dims"])
    src = {BIT: (z["bit_a"], z["bit_b"]), PROB: (z["prob_a"].view(np.uint8), z["prob_b"].view(np.uint8))}
    gv.set_option(gv.GVX_OPT_USE_SUMMARY, use_summary)
    try:
        for key, ta, tb, thr, off, cnt in z["collide_cases"]:
            ta, tb, thr, off, cnt = int(ta), int(tb), float(thr), int(off), int(cnt)
            a, b = _map(gv, ta, dims, src[ta][0]), _map(gv, tb, dims, src[tb][1])
            # linear offset -> Vector3i (the golden offsets are chosen to be representable)
            o = off
            oz, o = np.divmod(o, dims[0] * dims[1]) if o >= 0 else (0, o)
            oy, ox = np.divmod(o, dims[0]) if o >= 0 else (-1 if off == -dims[0] else 0, 0 if off == -dims[0] else o)
            assert (oz * dims[1] + oy) * dims[0] + ox == off
            got = a.collideWith(b, thr, (int(ox), int(oy), int(oz)))
            assert got == cnt, key
            assert np.array_equal(a.download().reshape(-1).view(np.uint8), z[key + "_marked"]), key
    finally:
        gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)


def test_collide_with_types_golden(gv):
    z = _load("colliders.npz")
    dims = tuple(int(v) for v in z["dims"])
    for thr in (1.0, 0.5, 0.3):
        a, b = _map(gv, BIT, dims, z["bit_a"]), _map(gv, PROB, dims, z["prob_b"].view(np.uint8))
        tv = gv.BitVectorVoxel()
        assert a.collideWithTypes(b, tv, thr) == int(z["types_%g_count" % thr][0])
        assert np.array_equal(tv.bytes, z["types_%g_mask" % thr])
        assert np.array_equal(a.download().reshape(-1), z["types_%g_marked" % thr])


def test_float_to_voxel_golden(gv):
    """insert one cloud, compare the set of occupied voxels with the reference's mapToVoxels of the same points"""
    z = _load("conversions.npz")
    for s in z["sides"]:
        pts, vox = z["pts_%g" % s], z["vox_%g" % s]
        dims = tuple(int(v) + 1 for v in vox.max(axis=0))
        m = gv.ProbVoxelMap(dims, float(s))
        m.insertPointCloud(pts, gv.eBVM_OCCUPIED)
        want = np.zeros(dims[0] * dims[1] * dims[2], bool)
        want[(vox[:, 2].astype(np.int64) * dims[1] + vox[:, 1]) * dims[0] + vox[:, 0]] = True
        assert np.array_equal(m.download() != -128, want), s


def test_sweptvolume_golden(gv):
    """every golden pair becomes one voxel of two BitVectorVoxelMaps; the window varies per pair, so pairs are grouped"""
    z = _load("sweptvolume.npz")
    v1, v2 = z["v1"], z["v2"]
    keys = sorted(set(zip(z["margin"].tolist(), z["sv_offset"].tolist())))
    checked = 0
    for margin, svo in keys:
        if (svo // 8) >= 32 + margin:
            continue  # shifts by >= 64 are undefined in the reference; refused by the library
        sel = np.flatnonzero((z["margin"] == margin) & (z["sv_offset"] == svo))
        dims = (len(sel), 1, 1)
        a, b = gv.BitVectorVoxelMap(dims, 1.0), gv.BitVectorVoxelMap(dims, 1.0)
        a.upload(v1[sel].copy()); b.upload(v2[sel].copy())
        tv = gv.BitVectorVoxel()
        cnt = a.collisionCheckBitvector(b, margin, tv, svo, mark=False)
        assert cnt == int(z["result_fresh"][sel].sum()), (margin, svo)
        assert np.array_equal(tv.bytes, np.bitwise_or.reduce(z["coll_fresh"][sel], axis=0)), (margin, svo)
        checked += len(sel)
    assert checked > 1000
    # shift: exact
    for sh in sorted(set(z["shift"].tolist())):
        sel = np.flatnonzero(z["shift"] == sh)
        m = gv.BitVectorVoxelMap((len(sel), 1, 1), 1.0)
        m.upload(v1[sel].copy())
        m.shiftLeftSweptVolumeIDs(sh)
        assert np.array_equal(m.download().reshape(-1, 32), z["shifted"][sel]), sh


def test_dh_chain_golden(gv):
    z = _load("dh_chain.npz")
    dh = np.concatenate([z["dh4"].astype(np.float64), z["joint_types"][:, None].astype(np.float64)], axis=1)
    got = gv.dh_chain_transforms(dh, z["joint_values"])
    assert np.array_equal(got.view(np.uint32), z["transforms_bits"])


def test_list_predicates_golden(gv):
    z = _load("list_predicates.npz")
    n = len(z["map_prob"])
    dims = (n, 1, 1)
    maps = {PROB: _map(gv, PROB, dims, z["map_prob"].view(np.uint8)), BIT: _map(gv, BIT, dims, z["map_bit"])}
    lst = gv.BitVectorVoxelList(dims, 1.0)
    ids = np.arange(n, dtype=np.uint32)
    for mt in (PROB, BIT):
        for thr in (1.0, 0.5, 30.0, -5.0):
            lst.upload(ids, None, z["list_voxels"].reshape(n, 32).copy())
            assert lst.collideWith(maps[mt], thr, mark=False) == int(z["hit_%d_%g_0" % (mt, thr)].sum())
            tv = gv.BitVectorVoxel()
            assert lst.collideWithTypes(maps[mt], tv, thr, mark=False) == int(z["hit_%d_%g_1" % (mt, thr)].sum())
            assert lst.collideWithTypeMask(maps[mt], z["mask"], thr, mark=False) == int(z["hit_%d_%g_2" % (mt, thr)].sum())
