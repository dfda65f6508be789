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
