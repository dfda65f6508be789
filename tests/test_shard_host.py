"""Host side of GVX_SHARD=z (gpu_voxels_b200/csrc/gvx_api_shard.cu), no GPU needed: the POSIX shared-memory rendezvous
and the host all-reduce (sum of counts, OR of masks) that carries every global collision result except the in-kernel
map x map count. World 3, one process per rank, through the C ABI with ctypes; plus the slab arithmetic the library and
sharding.py must agree on, and the facade program that runs on the GPU box must compile with a plain g++."""
import ctypes as C
import multiprocessing as mp
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
SHARD_SRC = os.path.join(ROOT, "tests", "cpp", "shard_example.cpp")
SHARD_EXE = os.path.join(ROOT, "tests", "cpp", "shard_example")


def build_shard_example():
    from _cppbuild import build_cpp, facade_headers
    return build_cpp(SHARD_SRC, SHARD_EXE, facade_headers())


def _rank_main(rank, world, name, rounds, q):
    try:
        from gpu_voxels_b200 import _lib
        L = _lib.load()
        _lib.check(L.gvx_shard_connect_host(rank, world, name.encode()))
        r_, w_, f_ = C.c_int(-1), C.c_int(-1), C.c_int(-1)
        L.gvx_shard_info(C.byref(r_), C.byref(w_), C.byref(f_))
        assert (r_.value, w_.value, f_.value) == (rank, world, 0)
        out = []
        for k in range(rounds):  # many rounds: both parities of the double buffer, ranks running ahead of each other
            v = np.zeros(1 + 256 + 4, np.uint64)
            v[0] = 1000 * k + rank
            v[1:257] = np.arange(256, dtype=np.uint64) * (rank + 1) + k
            v[257:] = [1 << (rank + 8 * (k % 5)), 0, (0xF0 << rank) & 0xFFFF, k if rank == 0 else 0]
            _lib.check(L.gvx_shard_allreduce_u64(v.ctypes.data_as(C.c_void_p), 257, 4))
            out.append(v.copy())
            if k % 3 == 0:  # an all-gather between the all-reduces (same buffers, same sequence): 64-byte "handles"
                mine = np.full(64, 16 * rank + (k & 15), np.uint8)
                mine[0] = k
                got = np.zeros((world, 64), np.uint8)
                _lib.check(L.gvx_shard_allgather(mine.ctypes.data_as(C.c_void_p), 64, got.ctypes.data_as(C.c_void_p)))
                for r in range(world):
                    assert got[r, 0] == k and (got[r, 1:] == 16 * r + (k & 15)).all(), (k, r, got[r, :4])
        one = np.array([rank == 1], np.uint64)   # an "any collision" flag: OR
        _lib.check(L.gvx_shard_allreduce_u64(one.ctypes.data_as(C.c_void_p), 0, 1))
        _lib.check(L.gvx_shard_disconnect())
        q.put((rank, out, int(one[0])))
    except Exception as e:  # pragma: no cover
        q.put((rank, repr(e), None))


def test_rendezvous_and_host_allreduce_world_3():
    world, rounds, name = 3, 40, "gvx_test_%d" % os.getpid()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_rank_main, args=(r, world, name, rounds, q)) for r in range(world)]
    for p in ps[1:]:   # ranks 1, 2 first: they must wait for rank 0's segment
        p.start()
    ps[0].start()
    res = sorted((q.get(timeout=120) for _ in range(world)), key=lambda t: t[0])
    for p in ps:
        p.join(timeout=30)
        assert p.exitcode == 0
    for rank, out, one in res:
        assert not isinstance(out, str), out
        assert one == 1
        for k, v in enumerate(out):
            assert v[0] == sum(1000 * k + r for r in range(world))
            assert np.array_equal(v[1:257], sum(np.arange(256, dtype=np.uint64) * (r + 1) + k for r in range(world)))
            want_or = [0, 0, 0, k]
            for r in range(world):
                want_or[0] |= 1 << (r + 8 * (k % 5))
                want_or[2] |= (0xF0 << r) & 0xFFFF
            assert v[257:].tolist() == want_or
    assert not os.path.exists("/dev/shm/" + name)   # rank 0 unlinked the segment


def test_single_process_is_a_no_op():
    from gpu_voxels_b200 import _lib
    L = _lib.load()
    v = np.array([5, 6], np.uint64)
    assert L.gvx_shard_allreduce_u64(v.ctypes.data_as(C.c_void_p), 1, 1) == 0
    assert v.tolist() == [5, 6]
    assert L.gvx_shard_connect_host(0, 1, b"unused") == 0
    assert L.gvx_shard_connect_host(3, 2, b"x") != 0 and b"bad rank" in L.gvx_last_error()


def test_library_slabs_equal_sharding_py():
    """gvx_map_create under GVX_SHARD=z cuts dz exactly like sharding.slab_for_rank (the Python-side partition)"""
    from gpu_voxels_b200 import sharding
    for dz in (1, 2, 7, 101, 150, 2048):
        for world in (1, 2, 3, 8, 16):
            if world > dz:
                continue
            base, rem = divmod(dz, world)
            for rank in range(world):
                z0 = rank * base + min(rank, rem)   # shard_slab(), gvx_api_shard.cu
                assert (z0, z0 + base + (1 if rank < rem else 0)) == sharding.slab_for_rank(dz, rank, world)


def test_shard_example_compiles_with_plain_gxx():
    assert os.path.exists(build_shard_example())


def test_halo_plan_enumerates_every_offset_pair_exactly_once():
    """gvx_halo_plan -- the arithmetic gvx_collide_map_map uses for an offset collide on z-slab maps (one launch per peer
    slab holding partners) -- against brute force: over all ranks and segments, the pairs (a index, b index) the launches
    visit must be exactly {(i + off, i) : 0 <= i < n, 0 <= i + off < n} (GV/voxelmap/TemplateVoxelMap.hpp:486-519: this
    map's pointer is shifted by the signed linear offset), each once, with the a index inside the visiting rank's slab
    and the b index inside rank s's slab."""
    from gpu_voxels_b200 import _lib, sharding
    L = _lib.load()
    rng = np.random.Generator(np.random.PCG64(3))
    cases = 0
    for dims in ((5, 4, 9), (3, 3, 7), (8, 2, 16), (1, 1, 5)):
        dx, dy, dz = dims
        layer, n = dx * dy, dx * dy * dz
        for world in (1, 2, 3, 4):
            if world > dz:
                continue
            even = sharding.slab_bounds(dz, world)
            cuts = np.sort(rng.choice(np.arange(0, dz + 1), size=world - 1, replace=True)).astype(np.uint32)
            ragged = np.concatenate([[0], cuts, [dz]]).astype(np.uint32)   # may hold EMPTY slabs
            for zb in (even, ragged):
                offs = [(0, 0, 1), (0, 0, -1), (1, 0, 0), (-1, 0, 0), (0, 1, 0), (dx - 1, dy - 1, dz - 1), (0, 0, dz), (0, 0, -dz - 3)]
                offs += [tuple(int(v) for v in rng.integers(-m, m + 1, 3)) for m in (1, 2, dz) for _ in range(4)]
                for off in offs:
                    lin = off[2] * layer + off[1] * dx + off[0]
                    want = {(i + lin, i) for i in range(n) if 0 <= i + lin < n}
                    got = []
                    for rank in range(world):
                        seg = np.zeros(3 * world, np.int64)
                        o3 = np.asarray(off, np.int32)
                        _lib.check(L.gvx_halo_plan(dx, dy, dz, zb.ctypes.data_as(C.c_void_p), world, rank,
                                                   o3.ctypes.data_as(C.c_void_p), seg.ctypes.data_as(C.c_void_p)))
                        a0, a1 = int(zb[rank]) * layer, int(zb[rank + 1]) * layer
                        for s in range(world):
                            b0, b1 = int(zb[s]) * layer, int(zb[s + 1]) * layer
                            i0, i1, ol = (int(v) for v in seg[3 * s:3 * s + 3])
                            assert 0 <= i0 <= i1 <= b1 - b0, (dims, world, zb, off, rank, s, i0, i1)
                            for il in range(i0, i1):
                                al = il + ol                       # what the kernel indexes: a_local[i' + off_local]
                                assert 0 <= al < a1 - a0, (dims, world, zb, off, rank, s, il, al)
                                got.append((a0 + al, b0 + il))
                    assert len(got) == len(set(got)) and set(got) == want, (dims, world, zb.tolist(), off, len(got), len(want))
                    cases += 1
    assert cases > 400
    bad = np.asarray([0, 5, 3, 9], np.uint32)   # decreasing bounds are refused
    seg = np.zeros(9, np.int64)
    o3 = np.zeros(3, np.int32)
    assert L.gvx_halo_plan(5, 4, 9, bad.ctypes.data_as(C.c_void_p), 3, 0, o3.ctypes.data_as(C.c_void_p), seg.ctypes.data_as(C.c_void_p)) != 0
