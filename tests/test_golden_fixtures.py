"""The CPU oracle against the golden vectors under tests/golden/, which were produced by the REFERENCE'S OWN code
compiled for the host (tests/golden/make_golden.py; oracle/ref_cpu_driver.cpp). Needs neither /root/reference nor
oracle/_ref nor a GPU: this is what pins the oracle wherever the tests run."""
import os

import numpy as np
import pytest

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
BIT, PROB = 0, 4


def _load(name):
    return np.load(os.path.join(G, name), allow_pickle=False)


def test_collide_map_map_all_type_pairs_thresholds_offsets(oracle):
    z = _load("colliders.npz")
    n = int(np.prod(z["dims"]))
    src = {BIT: (z["bit_a"], z["bit_b"]), PROB: (z["prob_a"].view(np.uint8), z["prob_b"].view(np.uint8))}
    assert len(z["collide_cases"]) == 64
    for key, ta, tb, thr, off, cnt in z["collide_cases"]:
        ta, tb, thr, off, cnt = int(ta), int(tb), float(thr), int(off), int(cnt)
        a = src[ta][0].copy()
        got = oracle.collide_map_map(a, ta, src[tb][1], tb, n, thr, off, True)
        assert got == cnt, key
        assert np.array_equal(a, z[key + "_marked"]), key


def test_collide_with_types(oracle):
    z = _load("colliders.npz")
    n = int(np.prod(z["dims"]))
    for thr in (1.0, 0.5, 0.3):
        a = z["bit_a"].copy()
        cnt, mask = oracle.collide_bit_prob_types(a, z["prob_b"], n, thr, True)
        assert cnt == int(z["types_%g_count" % thr][0])
        assert np.array_equal(mask, z["types_%g_mask" % thr]) and np.array_equal(a, z["types_%g_marked" % thr])


def test_conversions(oracle):
    z = _load("conversions.npz")
    assert [oracle.float_to_probability(float(t)) for t in z["thr"]] == list(z["f2p"])
    for s in z["sides"]:
        assert np.array_equal(oracle.map_to_voxels(float(s), z["pts_%g" % s]), z["vox_%g" % s]), s


def test_sweptvolume_window_and_shift(oracle):
    z = _load("sweptvolume.npz")
    for i in range(len(z["v1"])):
        c = z["coll_in"][i].copy()
        r, _ = oracle.bit_margin_check(z["v1"][i], z["v2"][i], int(z["margin"][i]), int(z["sv_offset"][i]), c)
        assert r == bool(z["result"][i]) and np.array_equal(c, z["coll_out"][i]), i
        assert np.array_equal(oracle.shift_left(z["v1"][i], int(z["shift"][i])), z["shifted"][i]), i


def test_dh_chain(oracle):
    z = _load("dh_chain.npz")
    got = oracle.dh_chain(z["dh4"], z["joint_types"], z["joint_values"])
    assert np.array_equal(got.view(np.uint32), z["transforms_bits"])


def test_list_entry_predicates(oracle):
    z = _load("list_predicates.npz")
    n = len(z["map_prob"])
    lst = oracle.VoxelList((n, 1, 1), 1.0)
    for mt, m in ((PROB, z["map_prob"].view(np.uint8)), (BIT, z["map_bit"])):
        for thr in (1.0, 0.5, 30.0, -5.0):
            for mode in (0, 1, 2):
                lst.ids = np.arange(n, dtype=np.uint32)
                lst.voxels = z["list_voxels"].reshape(n, 32).copy()
                cnt, _, _ = lst.collide(mode, m, mt, thr, 0, True, z["mask"])
                want = z["hit_%d_%g_%d" % (mt, thr, mode)]
                assert cnt == int(want.sum()), (mt, thr, mode)
                marked = (lst.voxels[:, 0] & 4) != 0
                was = (z["list_voxels"].reshape(n, 32)[:, 0] & 4) != 0
                assert np.array_equal(marked, was | want.astype(bool)), (mt, thr, mode)
