"""The header-only C++ facade (include/gpu_voxels/, the reference's class API over the C ABI): it must compile with a
plain g++ and link against libgvx.so (CPU check), and the reference's known-answer tests written against it must pass
on the GPU (tests/cpp/test_facade.cpp)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "test_facade.cpp")
EXE = os.path.join(ROOT, "tests", "cpp", "test_facade")
CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"


def _build():
    from _cppbuild import build_cpp, facade_headers
    build_cpp(SRC, EXE, [os.path.join(ROOT, "tests", "cpp", "mini_boost_test.h")] + facade_headers())


def test_facade_compiles_and_links_with_plain_gxx():
    _build()
    assert os.path.exists(EXE)


@pytest.mark.gpu
def test_reference_known_answer_tests_through_the_cpp_facade():
    _build()
    env = dict(os.environ, GVX_TEST_FILES=os.path.join(ROOT, "tests", "golden", "files"))
    r = subprocess.run([EXE, "!late_additions"], capture_output=True, text=True, timeout=600, env=env)
    print(r.stdout[-4000:], r.stderr[-4000:])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert " 0 failures" in r.stdout
