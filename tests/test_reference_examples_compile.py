"""Drop-in check of the C++ facade: example programs of the reference that stay on the path (dense maps, voxel lists,
collisions, shifting) are compiled UNMODIFIED from /root/reference against include/ and linked with libgvx.so by plain
g++. Nothing is copied: the sources are read where they lie, and the test is skipped where the reference is not mounted
(the GPU box). Examples that need what is out of scope (octrees: RobotVsEnvironment; Kinect / PCL:
SweptVolumeVsEnvironment; primitive arrays of the visualizer: Sandbox) are not in the list.

What makes them compile: the class API of include/gpu_voxels/, the std::-backed boost aliases of
helpers/BoostCompat.h, the icl_core logging / option stand-ins, and GpuVoxels::visualizeMap as a headless no-op."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXAMPLES = "/root/reference/packages/gpu_voxels/src/examples"
LINKAGE = "/root/reference/example_how_to_link"
CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"

pytestmark = pytest.mark.skipif(not os.path.isdir(EXAMPLES), reason="reference not mounted")


def _lib():
    from gpu_voxels_b200 import build as gvx_build
    gvx_build.build()
    return os.path.join(ROOT, "gpu_voxels_b200")


@pytest.mark.parametrize("name", ["Collisions", "VoxelMapSandbox", "VoxelListSandbox", "ShiftVsTransform", "Maps"])
def test_reference_example_builds_unmodified(name, tmp_path):
    exe = str(tmp_path / name)
    r = subprocess.run([CXX, "-std=c++14", "-O0", "-I" + os.path.join(ROOT, "include"), os.path.join(EXAMPLES, name + ".cpp"),
                        "-L" + _lib(), "-lgvx", "-Wl,-rpath," + _lib(), "-lpthread", "-o", exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    assert os.path.getsize(exe) > 10000


def test_reference_linkage_example_builds_unmodified(tmp_path):
    """example_how_to_link/: the library a downstream package builds on top of gpu_voxels and its main program"""
    exe = str(tmp_path / "gvl_linkage_test")
    r = subprocess.run([CXX, "-std=c++14", "-O0", "-I" + os.path.join(ROOT, "include"), "-I" + LINKAGE,
                        os.path.join(LINKAGE, "gvl_linkage_test_lib.cpp"), os.path.join(LINKAGE, "gvl_linkage_test.cpp"),
                        "-L" + _lib(), "-lgvx", "-Wl,-rpath," + _lib(), "-lpthread", "-o", exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
