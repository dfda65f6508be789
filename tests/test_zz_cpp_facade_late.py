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
