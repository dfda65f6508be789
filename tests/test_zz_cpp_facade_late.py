"""The facade cases written after the round's GPU minutes were spent (suite `late_additions` of
tests/cpp/test_facade.cpp: the GpuVoxelsMap base interface, list storage / dumps / collider entry points, the
reference's point-cloud equality and rotation tests). They compile and link in the CPU suite; this file sorts last so
that their first run on a GPU comes after every verified test."""
import os
import subprocess

import pytest

from test_cpp_facade import EXE, ROOT, _build


@pytest.mark.gpu
def test_late_facade_cases():
    _build()
    env = dict(os.environ, GVX_TEST_FILES=os.path.join(ROOT, "tests", "golden", "files"))
    r = subprocess.run([EXE, "late_additions"], capture_output=True, text=True, timeout=600, env=env)
    print(r.stdout[-4000:], r.stderr[-4000:])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert " 0 failures" in r.stdout and "late_additions/base_class_interface" in r.stdout


@pytest.mark.gpu
def test_map_info_reports_type_shape_slab_and_bytes():
    """gvx_map_info: what GpuVoxelsMap::getMapType / getDimensions / getMemoryUsage report, for both map types and a
    z-slab map"""
    import ctypes as C
    import gpu_voxels_b200 as gv
    from gpu_voxels_b200 import _lib
    L = _lib.init()
    cases = [(gv.ProbVoxelMap((40, 30, 20), 0.5), 4, (0, 20), 40 * 30 * 20),
             (gv.BitVectorVoxelMap((40, 30, 20), 0.25), 0, (0, 20), 40 * 30 * 20 * 32),
             (gv.ProbVoxelMap((40, 30, 20), 0.5, slab=(5, 12)), 4, (5, 12), 40 * 30 * 7)]
    for m, want_type, (z0, z1), want_bytes in cases:
        t, dims, side = C.c_int(-1), (C.c_uint32 * 3)(), C.c_float(0)
        a, b, nbytes = C.c_uint32(99), C.c_uint32(99), C.c_uint64(0)
        _lib.check(L.gvx_map_info(m.handle, C.byref(t), dims, C.byref(side), C.byref(a), C.byref(b), C.byref(nbytes)))
        assert (t.value, tuple(dims), a.value, b.value, nbytes.value) == (want_type, (40, 30, 20), z0, z1, want_bytes)
        assert side.value in (0.5, 0.25)
        _lib.check(L.gvx_map_info(m.handle, None, None, None, None, None, None))  # every output is optional
    assert L.gvx_map_info(C.c_uint64(1 << 60), None, None, None, None, None, None) == _lib.ERR_INVALID


@pytest.mark.gpu
def test_shutdown_releases_everything_and_init_works_again():
    """gvx_shutdown in a process of its own: maps, lists and buffers of the context are freed, calls fail with
    GVX_ERR_NOT_INITIALIZED until gvx_init runs again, a second shutdown is a no-op"""
    import sys
    code = r'''
import ctypes as C, numpy as np, sys
sys.path.insert(0, %r)
from gpu_voxels_b200 import _lib
L = _lib.init()
pts = np.random.Generator(np.random.PCG64(5)).uniform(0, 32, size=(5000, 3)).astype(np.float32)
pp = pts.ctypes.data_as(C.c_void_p)
def free_bytes():
    f = C.c_uint64(0); _lib.check(L.gvx_device_info(None, C.byref(f), None, None)); return f.value
free0 = free_bytes()
# raw handles on purpose: nothing on the Python side destroys them, gvx_shutdown has to
a, b, lst = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
_lib.check(L.gvx_map_create(C.byref(a), 4, 256, 256, 256, 1.0))     # 16 MiB
_lib.check(L.gvx_map_create(C.byref(b), 0, 256, 256, 128, 1.0))     # 256 MiB
_lib.check(L.gvx_list_create(C.byref(lst), 256, 256, 256, 1.0))
_lib.check(L.gvx_insert_points(a, pp, len(pts), _lib.MEM_HOST, 1, None))
_lib.check(L.gvx_list_insert_points(lst, pp, len(pts), _lib.MEM_HOST, 9))
free1 = free_bytes()
assert free0 - free1 > 250 << 20, (free0, free1)
assert L.gvx_shutdown() == 0 and L.gvx_shutdown() == 0
h, h2 = C.c_uint64(0), C.c_uint64(0)
assert L.gvx_map_create(C.byref(h), 4, 8, 8, 8, 1.0) == _lib.ERR_NOT_INITIALIZED
assert L.gvx_init(-1) == 0
assert free_bytes() - free1 > 250 << 20                              # the old context's memory came back
for m in (h, h2):
    assert L.gvx_map_create(C.byref(m), 4, 64, 64, 64, 1.0) == 0 and m.value != 0
    _lib.check(L.gvx_insert_points(m, pp, len(pts), _lib.MEM_HOST, 1, None))
n = C.c_uint64(0)
_lib.check(L.gvx_collide_map_map(h, h2, 0.5, None, 0, C.byref(n)))
assert n.value == len(np.unique(np.floor(pts).astype(np.int64) @ np.array([1, 64, 4096]))), n.value
assert L.gvx_shutdown() == 0
print("ok")
''' % ROOT
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok"), r.stdout[-2000:] + r.stderr[-3000:]


@pytest.mark.gpu
def test_list_file_with_a_hostile_entry_count_is_refused_without_allocating(tmp_path):
    """readFromDisk of a voxel list whose header announces 4 * 10^9 entries but carries two: GVX_ERR_IO, at once"""
    import struct
    import time
    import gpu_voxels_b200 as gv
    from gpu_voxels_b200 import _lib
    L = _lib.init()
    lst = gv.BitVectorVoxelList((64, 64, 64), 0.5)
    path = str(tmp_path / "hostile.lst")
    with open(path, "wb") as f:
        f.write(struct.pack("<i3IfI", 1, 64, 64, 64, 0.5, 4000000000) + b"\0" * 96)
    t0 = time.perf_counter()
    assert L.gvx_list_read_file(lst.handle, path.encode()) == _lib.ERR_IO
    assert time.perf_counter() - t0 < 1.0
    # and a well-formed file of two entries still loads
    with open(path, "wb") as f:
        f.write(struct.pack("<i3IfI", 1, 64, 64, 64, 0.5, 2) + struct.pack("<2I", 5, 70) +
                struct.pack("<6I", 5, 0, 0, 6, 1, 0) + (b"\x00\x08" + b"\0" * 30) * 2)
    _lib.check(L.gvx_list_read_file(lst.handle, path.encode()))
    n = _lib.C.c_uint64(0)
    _lib.check(L.gvx_list_size(lst.handle, _lib.C.byref(n)))
    assert n.value == 2
