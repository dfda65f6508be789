"""The C oracle against the reference's OWN voxel/collider code compiled for the host
(oracle/_ref/libgvref_cpu.so = GV/voxel/*.hpp, GV/helpers/BitVector.h, GV/voxelmap/kernels/VoxelMapOperations.h
included unmodified from /root/reference by oracle/ref_cpu_driver.cpp). Skipped where the .so was never built."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "oracle", "_ref", "libgvref_cpu.so")
pytestmark = pytest.mark.skipif(not os.path.exists(SO), reason="oracle/_ref/libgvref_cpu.so not built")
BIT, PROB = 0, 4


@pytest.fixture(scope="module")
def ref():
    L = C.CDLL(SO)
    L.gvrefcpu_collide_map_map.restype = C.c_uint64
    L.gvrefcpu_collide_map_map.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_uint64, C.c_float, C.c_int64, C.c_int]
    L.gvrefcpu_collide_bit_prob_types.restype = C.c_uint64
    L.gvrefcpu_collide_bit_prob_types.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_float, C.c_int, C.c_void_p]
    L.gvrefcpu_voxel_index_signed.restype = C.c_int64
    L.gvrefcpu_float_to_probability.argtypes = [C.c_float]
    L.gvrefcpu_insert_at.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_int]
    L.gvrefcpu_map_to_voxels.argtypes = [C.c_float, C.c_void_p, C.c_void_p]
    L.gvrefcpu_list_entry_hit.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_void_p]
    return L


def _random_maps(seed, n):
    rng = np.random.Generator(np.random.PCG64(seed))
    bit = np.zeros((n, 32), np.uint8)
    occ = rng.uniform(size=n) < 0.3
    for k in np.nonzero(occ)[0]:
        for b in rng.integers(0, 256, size=rng.integers(1, 4)):
            bit[k, b >> 3] |= 1 << (b & 7)
    bit[rng.uniform(size=n) < 0.1, 0] |= 1  # eBVM_FREE only / additionally
    prob = rng.integers(-128, 128, size=n).astype(np.int8)
    return bit.reshape(-1), prob


def test_float_to_probability(oracle, ref):
    for thr in np.linspace(0, 1, 101).astype(np.float32):
        assert oracle.float_to_probability(float(thr)) == ref.gvrefcpu_float_to_probability(float(thr))


def test_index_signed(oracle, ref):
    rng = np.random.Generator(np.random.PCG64(3))
    for _ in range(200):
        dims = rng.integers(1, 300, 3).astype(np.uint32)
        off = rng.integers(-50, 50, 3).astype(np.int32)
        assert oracle.voxel_index_signed(dims, off) == ref.gvrefcpu_voxel_index_signed(dims.ctypes.data_as(C.c_void_p), off.ctypes.data_as(C.c_void_p))


def test_map_to_voxels_nonnegative(oracle, ref):
    """host evaluation of the reference's mapToVoxels agrees with the oracle wherever it is defined (>= 0, finite)"""
    rng = np.random.Generator(np.random.PCG64(4))
    for side in (1.0, 0.01, 0.025, 0.013):
        p = rng.uniform(0, 300 * side, size=(3000, 3)).astype(np.float32)
        p[:100] = (rng.integers(0, 300, size=(100, 3)) * np.float32(side)).astype(np.float32)  # on voxel boundaries
        got = oracle.map_to_voxels(side, p)
        want = np.zeros_like(got)
        for i in range(len(p)):
            ref.gvrefcpu_map_to_voxels(side, p[i].ctypes.data_as(C.c_void_p), want[i].ctypes.data_as(C.c_void_p))
        assert np.array_equal(got, want)


def test_insert_values(oracle, ref):
    for t in (BIT, PROB):
        for mg in range(256):
            a = oracle.new_map(t, (1, 1, 1)); b = a.copy()
            oracle.insert_points(a, t, (1, 1, 1), 1.0, [[0.5, 0.5, 0.5]], mg)
            ref.gvrefcpu_insert_at(b.ctypes.data, 1 if t == PROB else 0, 0, mg)
            assert np.array_equal(a, b), (t, mg)
    for occ in range(-128, 128):
        for d in (72, -10, 127, -127, 1):
            m = np.array([occ], np.int8)
            oracle.lib().gvo_update_logodds  # noqa: B018  (the update rule is exercised through prob_update below)
            assert ref.gvrefcpu_update_occupancy(occ, d) == max(-127, min(127, occ + d))


@pytest.mark.parametrize("ta,tb", [(BIT, PROB), (PROB, BIT), (PROB, PROB), (BIT, BIT)])
def test_collide_map_map(oracle, ref, ta, tb):
    n = 20011
    bit1, prob1 = _random_maps(1, n)
    bit2, prob2 = _random_maps(2, n)
    for thr in (1.0, 0.7, 0.5, 0.1, 0.0):
        for off in (0, 1, -1, 177, -5000, 30000):
            a1 = (bit1 if ta == BIT else prob1).copy(); a2 = a1.copy()
            b = bit2 if tb == BIT else prob2
            c1 = oracle.collide_map_map(a1, ta, b, tb, n, thr, off, True)
            c2 = ref.gvrefcpu_collide_map_map(a2.ctypes.data, ta, b.ctypes.data, tb, n, thr, off, 1)
            assert c1 == c2, (thr, off)
            assert np.array_equal(a1, a2)


@pytest.mark.parametrize("ta,tb", [(BIT, PROB), (PROB, BIT), (PROB, PROB), (BIT, BIT)])
def test_collide_map_map_two_thresholds(oracle, ref, ta, tb):
    """DefaultCollider(Probability threshold1, Probability threshold2), GV/voxel/DefaultCollider.hpp:45-49"""
    ref.gvrefcpu_collide_map_map_thresholds.restype = C.c_uint64
    ref.gvrefcpu_collide_map_map_thresholds.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_uint64, C.c_int, C.c_int,
                                                        C.c_int64, C.c_int]
    n = 20011
    bit1, prob1 = _random_maps(3, n)
    bit2, prob2 = _random_maps(4, n)
    for t1, t2 in ((0, 0), (-128, 5), (5, -128), (-127, 127), (127, -127), (-20, 90), (64, -3), (-128, -128)):
        for off in (0, -1, 177):
            a1 = (bit1 if ta == BIT else prob1).copy(); a2 = a1.copy()
            b = bit2 if tb == BIT else prob2
            c1 = oracle.collide_map_map_thresholds(a1, ta, b, tb, n, t1, t2, off, True)
            c2 = ref.gvrefcpu_collide_map_map_thresholds(a2.ctypes.data, ta, b.ctypes.data, tb, n, t1, t2, off, 1)
            assert c1 == c2, (t1, t2, off)
            assert np.array_equal(a1, a2)
    # the one-threshold form is the two-threshold form with t1 == t2 == floatToProbability(thr)
    a1 = prob1.copy()
    assert oracle.collide_map_map(a1, PROB, prob2, PROB, n, 0.5, 0, False) == \
        oracle.collide_map_map_thresholds(a1, PROB, prob2, PROB, n, 0, 0, 0, False)


def test_collide_types(oracle, ref):
    n = 20011
    bit1, _ = _random_maps(5, n)
    _, prob2 = _random_maps(6, n)
    for thr in (1.0, 0.5, 0.2):
        a1, a2 = bit1.copy(), bit1.copy()
        m1 = np.zeros(32, np.uint8); m1[31] = 0x80; m2 = m1.copy()
        c1, m1 = oracle.collide_bit_prob_types(a1, prob2, n, thr, True, m1)
        c2 = ref.gvrefcpu_collide_bit_prob_types(a2.ctypes.data, prob2.ctypes.data, n, thr, 1, m2.ctypes.data)
        assert c1 == c2 and np.array_equal(m1, m2) and np.array_equal(a1, a2)


def test_list_entry_predicates(oracle, ref):
    """mode 0/1/2 hit predicates of the list x map kernels, entry by entry"""
    n = 4000
    bitm, probm = _random_maps(7, n)
    bitl, _ = _random_maps(8, n)
    mask = np.zeros(32, np.uint8); mask[4] = 0xFF; mask[20] = 0x0F
    lst = oracle.VoxelList((n, 1, 1), 1.0)
    for mt, m in ((PROB, probm), (BIT, bitm)):
        for thr in (1.0, 0.5, 30.0, -5.0):
            for mode in (0, 1, 2):
                lst.ids = np.arange(n, dtype=np.uint32); lst.voxels = bitl.reshape(n, 32).copy()
                cnt, _, _ = lst.collide(mode, m, mt, thr, 0, False, mask)
                vb = 1 if mt == PROB else 32
                want = sum(ref.gvrefcpu_list_entry_hit(mode, bitl[32 * i:].ctypes.data, m[vb * i:].ctypes.data, mt, thr, mask.ctypes.data) for i in range(n))
                assert cnt == want, (mt, thr, mode)


def test_dh_chain_against_reference_kinematic_link(oracle, ref):
    """DHParameters::convertDHtoM + KinematicChain's accumulation: the reference's own KinematicLink.cu and
    Matrix4f operator*, compiled for the host, against the restatement -- bit for bit, incl. huge / tiny / -0 angles"""
    rng = np.random.Generator(np.random.PCG64(21))
    nl, nc = 9, 30000
    dh = rng.uniform(-1, 1, size=(nl, 4)).astype(np.float32)
    dh[:, 3] = rng.choice(np.array([0, np.pi / 2, -np.pi / 2, np.pi, 1.1], np.float32), nl)
    jt = rng.integers(0, 2, nl).astype(np.int32)
    jv = rng.uniform(-7, 7, size=(nc, nl)).astype(np.float32)
    jv[:8, :] = np.array([0.0, -0.0, np.pi / 2, 1e-30, 1e5, -12345.678, 3e38, 1e-45], np.float32)[:, None]
    got = oracle.dh_chain(dh, jt, jv)
    want = np.empty_like(got)
    ref.gvrefcpu_dh_chain(dh.ctypes.data_as(C.c_void_p), jt.ctypes.data_as(C.c_void_p), nl,
                          jv.ctypes.data_as(C.c_void_p), nc, want.ctypes.data_as(C.c_void_p))
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_bit_margin_check_and_shift_against_reference_bitvector_h(oracle, ref):
    """bitMarginCollisionCheck / performLeftShift (GV/helpers/BitVector.h:361-471) on random vectors, windows and
    offsets -- including the result vector's untouched bytes (it is not reset by the reference)"""
    rng = np.random.Generator(np.random.PCG64(31))
    for it in range(20000):
        dens = rng.choice([0.02, 0.1, 0.5])
        v1 = np.packbits(rng.uniform(size=256) < dens, bitorder="little")
        v2 = np.packbits(rng.uniform(size=256) < dens, bitorder="little")
        margin = int(rng.integers(0, 32))
        svo = int(rng.integers(0, 8 * (32 + margin))) if it % 3 == 0 else int(rng.integers(0, 24))
        c1 = rng.integers(0, 256, 32).astype(np.uint8)
        c2 = c1.copy()
        r1, _ = oracle.bit_margin_check(v1, v2, margin, svo, c1)
        r2 = ref.gvrefcpu_bit_margin_check(v1.ctypes.data_as(C.c_void_p), v2.ctypes.data_as(C.c_void_p),
                                           c2.ctypes.data_as(C.c_void_p), margin, svo)
        assert r1 == bool(r2) and np.array_equal(c1, c2), (margin, svo)
        sh = int(rng.integers(0, 57))
        s2 = v1.copy()
        ref.gvrefcpu_shift_left(s2.ctypes.data_as(C.c_void_p), sh)
        assert np.array_equal(oracle.shift_left(v1, sh), s2), sh
