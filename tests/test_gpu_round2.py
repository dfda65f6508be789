"""GPU parity of the round-2 fixes, through the C ABI, against the CPU oracle (==, no tolerance):
  * thresholds that convert to int8 -128 / raw float thresholds <= -128: UNKNOWN voxels collide in the reference
    (GV/voxel/DefaultCollider.hpp:51-98, GV/voxel/ProbabilisticVoxel.hpp:94-98); the "known" summary must be bypassed
  * Prob x Prob ranges shorter than one 16-byte vector (summaries off / offsets)
  * early-exit boolean collisionCheck (kernelCollideVoxelMaps, GV/voxelmap/kernels/VoxelMapOperations.hpp:78-120)
  * raw device pointer: sticky invalidation, const accessor, fallback counter
  * GVX_MEM_DEVICE clouds produced on a torch stream; page-locked GVX_MEM_HOST clouds reused right after the call
"""
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


def _mk(gv, t, dims, side, slab=None):
    return (gv.ProbVoxelMap if t == PROB else gv.BitVectorVoxelMap)(dims, side, slab=slab)


def _cloud(seed, n, hi):
    return np.random.Generator(np.random.PCG64(seed)).uniform(0.0, hi, size=(n, 3)).astype(np.float32)


# ---- thresholds at which UNKNOWN voxels collide ----------------------------------------------------------------------
@pytest.mark.parametrize("thr", [-0.005, -1.012, 0.0, 0.5])
@pytest.mark.parametrize("pair", [(PROB, BIT), (BIT, PROB), (PROB, PROB)])
def test_map_map_threshold_minus_128(gv, oracle, pair, thr):
    """float_to_probability(-0.005) == float_to_probability(-1.012) == -128: `occ >= -128` holds for UNKNOWN voxels, so
    every occupied voxel of the other side collides. Summary path on and off must both equal the oracle."""
    ti = oracle.float_to_probability(thr)
    if thr in (-0.005, -1.012):
        assert ti == -128
    dims, side = (67, 45, 39), 0.01   # 117 585 voxels: a partial last group and a partial last quad
    n = dims[0] * dims[1] * dims[2]
    pts = [_cloud(11, 9000, 0.39), _cloud(12, 7000, 0.39)]
    for use_summary in (1, 0):
        gv.set_option(gv.GVX_OPT_USE_SUMMARY, use_summary)
        try:
            maps, refs = [], []
            for t, p in zip(pair, pts):
                m, r = _mk(gv, t, dims, side), oracle.new_map(t, dims)
                if t == PROB:
                    pose = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0.001]], np.float32)
                    m.insertSensorData(p, pose, delta_occupied=-20)   # known but unoccupied at thr 0.5
                    oracle.update_logodds(r, dims, side, p, pose, -20)
                    m.insertPointCloud(p[:3000], 1)
                    oracle.insert_points(r, t, dims, side, p[:3000], 1)
                else:
                    m.insertPointCloud(p, 9)
                    oracle.insert_points(r, t, dims, side, p, 9)
                maps.append(m); refs.append(r)
            for mark in (False, True):
                want = oracle.collide_map_map(refs[0], pair[0], refs[1], pair[1], n, thr, 0, mark)
                got = maps[0].collideWith(maps[1], thr, mark=mark)
                assert got == want, (pair, thr, use_summary, mark, got, want)
                assert np.array_equal(maps[0].download().reshape(-1).view(np.uint8), refs[0].view(np.uint8))
                # the marks may have turned UNKNOWN voxels into known ones: the summary must have followed
                bits, valid = maps[0].summary()
                if valid and pair[0] == PROB:
                    assert np.array_equal(bits, maps[0].download() != -128), (pair, thr, use_summary, mark)
            if use_summary and pair[0] == PROB:   # and the summary path after the marks agrees with the oracle again
                want2 = oracle.collide_map_map(refs[0].copy(), pair[0], refs[1], pair[1], n, 0.5, 0, False)
                assert maps[0].collideWith(maps[1], 0.5, mark=False) == want2
            if ti == -128 and pair == (PROB, BIT):  # every occupied robot voxel collides, known environment or not
                v = refs[1].reshape(-1, 32)
                assert want == int((((v[:, 0] & 0xFE) != 0) | v[:, 1:].any(1)).sum())
        finally:
            gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)


@pytest.mark.parametrize("thr", [-128.0, -200.0, -127.5, 1.0])
@pytest.mark.parametrize("mode", ["types", "mask"])
def test_list_map_raw_threshold_below_minus_128(gv, oracle, mode, thr):
    """collideWithTypes / collideWithTypeMask test a ProbabilisticVoxel with the RAW float: (float)occ >= thr is true for
    UNKNOWN (-128) when thr <= -128 -- every in-range list entry collides, known or not."""
    dims, side = (40, 33, 29), 0.01
    pl, pm = _cloud(21, 6000, 0.29), _cloud(22, 5000, 0.29)
    for use_summary in (1, 0):
        gv.set_option(gv.GVX_OPT_USE_SUMMARY, use_summary)
        try:
            lst = gv.BitVectorVoxelList(dims, side)
            lst.insertMetaPointCloud(gv.MetaPointCloud([pl[:3000], pl[3000:]]), [11, 12])
            ref = oracle.VoxelList(dims, side)
            ref.insert_meta(pl, [3000, 3000], [11, 12])
            m, r = gv.ProbVoxelMap(dims, side), oracle.new_map(PROB, dims)
            m.insertPointCloud(pm, 1)
            oracle.insert_points(r, PROB, dims, side, pm, 1)
            if mode == "types":
                want, wmask, _ = ref.collide(1, r, PROB, thr, 0, False)
                types = gv.BitVectorVoxel()
                got = lst.collideWithTypes(m, types, thr, None, mark=False)
                assert got == want and np.array_equal(types.bytes, wmask), (thr, use_summary, got, want)
            else:
                mk = gv.BitVectorVoxel(); mk.setBit(12)
                want = ref.collide(2, r, PROB, thr, 0, False, mask_in=mk.bytes)[0]
                got = lst.collideWithTypeMask(m, mk, thr, None, mark=False)
                assert got == want, (thr, use_summary, got, want)
            if thr <= -128.0 and mode == "types":
                assert want == len(ref)
        finally:
            gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)


@pytest.mark.parametrize("thr", [-0.005, 0.5])
def test_list_map_default_collider_minus_128(gv, oracle, thr):
    """collideWith / per-meaning counts use DefaultCollider's int8 threshold: -128 lets UNKNOWN map voxels collide"""
    dims, side = (40, 33, 29), 0.01
    pl, pm = _cloud(23, 6000, 0.29), _cloud(24, 5000, 0.29)
    lst = gv.BitVectorVoxelList(dims, side)
    lst.insertMetaPointCloud(gv.MetaPointCloud([pl[:3000], pl[3000:]]), [40, 41])
    ref = oracle.VoxelList(dims, side)
    ref.insert_meta(pl, [3000, 3000], [40, 41])
    m, r = gv.ProbVoxelMap(dims, side), oracle.new_map(PROB, dims)
    m.insertPointCloud(pm, 1)
    oracle.insert_points(r, PROB, dims, side, pm, 1)
    want = ref.collide(0, r, PROB, thr, 0, False)[0]
    assert lst.collideWith(m, thr, None, mark=False) == want
    w_tot, _, w_per = ref.collide(3, r, PROB, thr, 0, False)
    tot, per = lst.collideCountingPerMeaning(m, None, thr, None, mark=False)
    assert tot == w_tot and np.array_equal(per, w_per)
    if thr == -0.005:
        assert want == len(ref)


# ---- Prob x Prob streaming kernel on ranges shorter than a 16-byte vector --------------------------------------------
def test_prob_prob_short_ranges_without_summary(gv, oracle):
    """with the 16-aligned middle empty (vb == ve) the tail [vb, i_end) was dropped"""
    gv.set_option(gv.GVX_OPT_USE_SUMMARY, 0)
    try:
        for n in list(range(1, 41)) + [47, 48, 49, 63, 64, 65]:
            dims, side = (n, 1, 1), 1.0
            rng = np.random.Generator(np.random.PCG64(100 + n))
            va = rng.choice(np.array([-128, -127, -5, 0, 3, 127], np.int8), size=n)
            vb = rng.choice(np.array([-128, -127, -5, 0, 3, 127], np.int8), size=n)
            a, b = gv.ProbVoxelMap(dims, side), gv.ProbVoxelMap(dims, side)
            a.upload(va); b.upload(vb)
            for off in (0, 1, -1, 5, -7, 16, -16, 17):
                if abs(off) >= n:
                    continue
                ra = va.copy()
                want = oracle.collide_map_map(ra, PROB, vb, PROB, n, 0.5, off, True)
                a.upload(va)
                got = a.collideWith(b, 0.5, (off, 0, 0), mark=True)
                assert got == want, (n, off, got, want)
                assert np.array_equal(a.download(), ra), (n, off)
    finally:
        gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)


# ---- boolean collisionCheck with early exit ---------------------------------------------------------------------------
@pytest.mark.parametrize("pair", [(PROB, BIT), (BIT, PROB), (BIT, BIT), (PROB, PROB)])
def test_collision_check_early_exit_both_outcomes(gv, oracle, pair):
    """flag == (count != 0) on every path (summaries on / off), no write-back, and the early-exit flag is lowered
    again: a counting call right after an early exit still returns the full count."""
    dims, side = (96, 64, 80), 0.01
    n = dims[0] * dims[1] * dims[2]
    pa, pb = _cloud(31, 40000, 0.6), _cloud(32, 40000, 0.6)
    pb_far = pb + np.float32(0.0)
    pb_far[:, 0] = pb_far[:, 0] * np.float32(0.3) + np.float32(0.7)   # x >= 0.7: beyond pa's x < 0.6
    for use_summary in (1, 0):
        gv.set_option(gv.GVX_OPT_USE_SUMMARY, use_summary)
        try:
            for cloud_b, expect in ((pb, True), (pb_far, False)):
                a, b = _mk(gv, pair[0], dims, side), _mk(gv, pair[1], dims, side)
                ra, rb = oracle.new_map(pair[0], dims), oracle.new_map(pair[1], dims)
                a.insertPointCloud(pa, 1); oracle.insert_points(ra, pair[0], dims, side, pa, 1)
                b.insertPointCloud(cloud_b, 1); oracle.insert_points(rb, pair[1], dims, side, cloud_b, 1)
                want = oracle.collide_map_map(ra.copy(), pair[0], rb, pair[1], n, 1.0, 0, False)
                assert (want != 0) == expect
                for _ in range(3):
                    assert a.collisionCheck(b, 1.0) == expect
                    assert a.collideWith(b, 1.0, mark=False) == want   # the flag of the early exit does not leak
                assert np.array_equal(a.download().reshape(-1).view(np.uint8), ra.view(np.uint8))  # no write-back
        finally:
            gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)


def test_collision_check_early_exit_is_early(gv):
    """a colliding 512^3 BitVector x BitVector pair, voxels streamed (8.6 GB, ~1.2 ms): the boolean check leaves after
    the first iterations instead of reading both grids (device time, CUDA events on a stream shared with the library)"""
    import torch
    L = gv._lib.load()
    dims, side = (512, 512, 512), 0.01
    a, b = gv.BitVectorVoxelMap(dims, side), gv.BitVectorVoxelMap(dims, side)
    p = _cloud(33, 200000, 5.12)
    a.insertPointCloud(p, 1); b.insertPointCloud(p, 1)
    stream = torch.cuda.Stream()
    gv._lib.set_stream(stream.cuda_stream)
    gv.set_option(gv.GVX_OPT_USE_SUMMARY, 0)
    try:
        def timed(fn, reps=5):
            best = 1e9
            for _ in range(reps):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                fn()
                e1.record(stream)
                torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1))
            return best
        full = timed(lambda: a.collideWith(b, 1.0, mark=False))
        early = timed(lambda: a.collisionCheck(b, 1.0))
        assert a.collisionCheck(b, 1.0) is True
        assert full > 0.8, full                       # 8.6 GB cannot be read in less
        assert early < 0.1 * full, (early, full)
    finally:
        gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)
        gv._lib.set_stream(None)
        a.destroy(); b.destroy()


# ---- raw device pointer: sticky invalidation, const accessor ---------------------------------------------------------
def test_raw_pointer_keeps_summary_invalid_until_refresh(gv, oracle):
    L = gv._lib.load()
    dims, side = (64, 48, 40), 0.01
    n = dims[0] * dims[1] * dims[2]
    m, other = gv.ProbVoxelMap(dims, side), gv.BitVectorVoxelMap(dims, side)
    pts = _cloud(41, 20000, 0.4)
    m.insertPointCloud(pts, 1)
    other.insertPointCloud(pts, 5)
    want = m.collideWith(other, 1.0, mark=False)
    assert want > 0
    # read-only accessor: the summary stays valid
    p = C.c_void_p()
    gv._lib.check(L.gvx_map_device_ptr_const(m.handle, C.byref(p), None, None))
    assert p.value and m.summary()[1]
    f0 = L.gvx_summary_fallback_count()
    ptr = m.getVoidDeviceDataPtr()
    assert ptr == p.value and not m.summary()[1]
    m.clearMap()                      # does NOT re-validate: the caller still holds the writable pointer
    assert not m.summary()[1]
    # the caller writes voxels behind the library's back, AFTER the clear
    raw = np.full(n, -128, np.int8)
    raw[::3] = 127
    alias = gv.ProbVoxelMap(dims, side, _wrap_ptr=ptr)   # the caller's own view of the same memory (no summary)
    alias.upload(raw)
    alias.destroy()
    ro = oracle.new_map(BIT, dims); oracle.insert_points(ro, BIT, dims, side, pts, 5)
    w2 = oracle.collide_map_map(raw.copy(), PROB, ro, BIT, n, 1.0, 0, False)
    assert w2 > 0
    assert m.collideWith(other, 1.0, mark=False) == w2      # streams the voxels: sees the direct writes
    assert L.gvx_summary_fallback_count() > f0
    m.refreshSummary()
    assert m.summary()[1]
    assert m.collideWith(other, 1.0, mark=False) == w2
    m.clearMap()                      # pointer no longer "out": clear re-validates as before
    assert m.summary()[1] and m.collideWith(other, 1.0, mark=False) == 0


# ---- stream contracts ---------------------------------------------------------------------------------------------------
def test_device_cloud_produced_by_a_long_torch_kernel(gv, oracle):
    """libgvx enqueues on its own non-blocking stream; a CUDA tensor still being produced on torch's stream must
    not be read early. The producer here takes milliseconds (a chain of large matmuls feeding the cloud)."""
    import torch
    dims, side = (128, 128, 64), 0.01
    base = torch.from_numpy(_cloud(51, 300000, 1.28)).cuda()
    torch.cuda.synchronize()
    want = oracle.new_map(PROB, dims)
    oracle.insert_points(want, PROB, dims, side, base.cpu().numpy(), 1)
    for rep in range(3):
        big = torch.randn(4096, 4096, device="cuda")
        for _ in range(12):               # a few ms of queued work ahead of the cloud
            big = (big @ big).clamp_(-1, 1)
        pts = (base + big[0, 0] * 0.0).contiguous()   # depends on the matmul chain; values == base (x*0 == 0, finite)
        m = gv.ProbVoxelMap(dims, side)
        m.insertPointCloud(pts, 1)        # CUDA tensor -> GVX_MEM_DEVICE
        assert np.array_equal(m.download(), want), "cloud was read before torch had produced it (rep %d)" % rep


def test_pinned_host_cloud_may_be_reused_when_the_call_returns(gv, oracle):
    """GVX_MEM_HOST: 'consumed or staged when the call returns' also for page-locked memory"""
    import torch
    dims, side = (128, 128, 64), 0.01
    src = _cloud(52, 400000, 1.28)
    want = oracle.new_map(BIT, dims)
    oracle.insert_points(want, BIT, dims, side, src, 7)
    pinned = torch.from_numpy(src.copy()).pin_memory()
    for rep in range(4):
        pinned.copy_(torch.from_numpy(src))
        m = gv.BitVectorVoxelMap(dims, side)
        m.insertPointCloud(pinned.numpy(), 7, check_outside=False)   # numpy view of pinned memory -> GVX_MEM_HOST
        pinned.fill_(float("nan"))                                   # the caller reuses its buffer at once
        assert np.array_equal(m.download().reshape(-1), want)


# ---- the multi-frame fold's division-free slab test -----------------------------------------------------------------
@pytest.mark.parametrize("side", [0.01, 0.025, 0.3, 1.0, 0.007])
def test_fold_slab_filter_is_exact_at_layer_boundaries(gv, oracle, side):
    """k_update_logodds_frames drops the points of other z-slabs with lo <= wz < hi (two floats found on the host by
    stepping ulps, gvx_insert.cu fold_z_bound) instead of floor(wz / side): by monotonicity that is the same predicate.
    Points a few ulps around EVERY layer boundary, slabs cut at awkward places: each slab must be byte-identical to its
    slice of the oracle's unsharded map."""
    dims = (8, 6, 37)
    s32 = np.float32(side)
    zs = []
    for layer in range(0, dims[2] + 2):
        w = np.float32(layer) * s32
        cand = [w]
        for _ in range(4):
            cand.append(np.nextafter(cand[-1], np.float32(np.inf), dtype=np.float32))
        lo = w
        for _ in range(4):
            lo = np.nextafter(lo, np.float32(-np.inf), dtype=np.float32)
            cand.append(lo)
        zs += cand
    zs = np.asarray(zs, np.float32)
    rng = np.random.Generator(np.random.PCG64(3))
    pts = np.stack([rng.uniform(0, dims[0] * side, len(zs)), rng.uniform(0, dims[1] * side, len(zs)), zs], -1).astype(np.float32)
    pts = np.concatenate([pts, rng.uniform(0, 1, size=(5000, 3)).astype(np.float32) * (np.asarray(dims) * side).astype(np.float32)])
    pts[7] = np.nan
    pose = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0]], np.float32)   # world z == sensor z: bit-for-bit the values above
    want = oracle.new_map(PROB, dims)
    frames = [pts[: len(pts) // 2], pts[len(pts) // 2:]]
    for f in frames:
        oracle.update_logodds(want, dims, side, f, pose)
    plane = dims[0] * dims[1]
    for cuts in ([0, 37], [0, 1, 2, 36, 37], [0, 5, 11, 12, 20, 29, 37]):
        for z0, z1 in zip(cuts[:-1], cuts[1:]):
            m = gv.ProbVoxelMap(dims, side, slab=(z0, z1))
            m.insertSensorDataMulti(frames, [pose, pose])
            assert np.array_equal(m.download(), want[z0 * plane:z1 * plane]), (side, z0, z1)
            bits, valid = m.summary()
            assert valid and np.array_equal(bits, want[z0 * plane:z1 * plane] != -128)


# ---- DefaultCollider(Probability threshold1, Probability threshold2) -----------------------------------------------------
@pytest.mark.parametrize("pair", [(PROB, BIT), (BIT, PROB), (PROB, PROB), (BIT, BIT)])
def test_two_threshold_collider(gv, oracle, pair):
    """VERDICT r1 missing #7 (GV/voxel/DefaultCollider.hpp:45-49): threshold1 tests side a's ProbabilisticVoxels, threshold2
    side b's; incl. -128 on one side only (that side's 'known' summary must be bypassed, the other side's stays in use),
    offsets, marks, summaries on and off."""
    dims, side = (67, 45, 39), 0.01
    n = dims[0] * dims[1] * dims[2]
    rng = np.random.Generator(np.random.PCG64(77))
    vals = np.array([-128, -128, -128, -127, -60, -1, 0, 1, 40, 127], np.int8)
    prob_a, prob_b = rng.choice(vals, size=n), rng.choice(vals, size=n)
    bit_a = np.zeros((n, 32), np.uint8); bit_b = np.zeros((n, 32), np.uint8)
    bit_a[rng.uniform(size=n) < 0.2, 5] = 2; bit_b[rng.uniform(size=n) < 0.2, 9] = 1
    bit_a[rng.uniform(size=n) < 0.1, 0] |= 1; bit_b[rng.uniform(size=n) < 0.1, 0] |= 1   # eBVM_FREE does not occupy
    src_a, src_b = (prob_a if pair[0] == PROB else bit_a), (prob_b if pair[1] == PROB else bit_b)
    a, b = _mk(gv, pair[0], dims, side), _mk(gv, pair[1], dims, side)
    b.upload(src_b)
    for use_summary in (1, 0):
        gv.set_option(gv.GVX_OPT_USE_SUMMARY, use_summary)
        try:
            for t1, t2 in ((0, 0), (-128, 5), (5, -128), (-127, 127), (127, -127), (-20, 90), (64, -3), (-128, -128)):
                for off, mark in (((0, 0, 0), False), ((0, 0, 0), True), ((3, -1, 2), True)):
                    a.upload(src_a)
                    ra = src_a.copy().reshape(-1)
                    lin = off[2] * dims[0] * dims[1] + off[1] * dims[0] + off[0]
                    want = oracle.collide_map_map_thresholds(ra, pair[0], src_b.reshape(-1), pair[1], n, t1, t2, lin, mark)
                    for rep in range(2):   # the second call also goes through the coarse mask
                        if rep:
                            want = oracle.collide_map_map_thresholds(ra, pair[0], src_b.reshape(-1), pair[1], n, t1, t2, lin, mark)
                        got = a.collisionCheckWithCounter(b, t1, t2, off, mark=mark)
                        assert got == want, (pair, t1, t2, off, mark, use_summary, rep, got, want)
                    assert np.array_equal(a.download().reshape(-1).view(np.uint8), ra.view(np.uint8))
        finally:
            gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)
    with pytest.raises(gv.GvxError):
        a.collisionCheckWithCounter(b, 200, 0)
