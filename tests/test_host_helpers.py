"""Host-side helpers of the C++ facade that need no GPU: computeLinearLoad / interpolateLinear
(include/gpu_voxels/helpers/MathHelpers.h; SURVEY 8 row a9, GV/helpers/MathHelpers.cpp:29-160). A small C++ program
built with plain g++ prints the facade's results; they must equal tests/golden/math_helpers.npz bit for bit -- the
output of the reference's own MathHelpers.cpp compiled for the host (tests/golden/make_golden_files.py)."""
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "cpp", "host_helpers.cpp")
EXE = os.path.join(ROOT, "tests", "cpp", "host_helpers")
CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"


def _exe():
    hdr = os.path.join(ROOT, "include", "gpu_voxels", "helpers", "MathHelpers.h")
    if not (os.path.exists(EXE) and os.path.getmtime(EXE) >= max(os.path.getmtime(SRC), os.path.getmtime(hdr))):
        subprocess.check_call([CXX, "-std=c++14", "-O2", "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(ROOT, "include"),
                               SRC, "-o", EXE])
    return EXE


def _run(mode, args):
    out = []
    args = [str(int(a)) for a in args]
    for i in range(0, len(args), 600):  # keep the command line short
        r = subprocess.run([_exe(), mode] + args[i:i + 600], capture_output=True, text=True, check=True)
        out += r.stdout.split("\n")[:-1]
    return out


def test_compute_linear_load_equals_reference():
    g = np.load(os.path.join(ROOT, "tests", "golden", "math_helpers.npz"))
    lines = _run("load", g["n"])
    got = np.array([[int(v) for v in ln.split()] for ln in lines], np.uint32)
    assert np.array_equal(got, g["load"])
    # the properties callers rely on: enough threads for every item up to the grid limit, never an empty launch
    n = g["n"].astype(np.uint64)
    small = n <= 65535 * 1024
    assert (got[small, 0].astype(np.uint64) * got[small, 1] >= n[small]).all() and (got >= 1).all() and got[:, 0].max() == 65535


def test_interpolate_linear_equals_reference():
    g = np.load(os.path.join(ROOT, "tests", "golden", "math_helpers.npz"))
    got = np.array([int(v) for v in _run("lerpf", g["f"].view(np.uint32).reshape(-1))], np.uint32)
    assert np.array_equal(got, g["lerp_f"].view(np.uint32))
    gotd = np.array([int(v) for v in _run("lerpd", g["d"].view(np.uint64).reshape(-1))], np.uint64)
    assert np.array_equal(gotd, g["lerp_d"].view(np.uint64))
    assert subprocess.run([_exe(), "vec"], capture_output=True, text=True, check=True).stdout.strip() == "ok"


def test_probabilistic_voxel_value_type_equals_reference():
    """include/gpu_voxels/voxel/ProbabilisticVoxel.h against the reference's header for EVERY (occupancy, delta) pair and
    every meaning (tests/golden/prob_voxel.npz, from GV/voxel/ProbabilisticVoxel.hpp compiled for the host)"""
    g = np.load(os.path.join(ROOT, "tests", "golden", "prob_voxel.npz"))
    hdr = os.path.join(ROOT, "include", "gpu_voxels", "voxel", "ProbabilisticVoxel.h")
    if os.path.exists(EXE) and os.path.getmtime(EXE) < os.path.getmtime(hdr):
        os.remove(EXE)
    lines = subprocess.run([_exe(), "prob"], capture_output=True, text=True, check=True).stdout.split("\n")
    upd = np.array([[int(v) for v in ln.split()] for ln in lines[:256]], np.int8)
    assert np.array_equal(upd, g["update"])
    assert np.array_equal(np.array([int(v) for v in lines[256].split()], np.int8), g["insert"])
    assert lines[257] == "ok"


def test_bitvector_window_check_and_shift_equal_reference():
    """bitMarginCollisionCheck / performLeftShift of the facade (include/gpu_voxels/voxel/BitVoxel.h) on the 1500 golden
    vectors recorded from the reference's BitVector.h (tests/golden/sweptvolume.npz): results and every output byte"""
    g = np.load(os.path.join(ROOT, "tests", "golden", "sweptvolume.npz"))
    hdr = os.path.join(ROOT, "include", "gpu_voxels", "voxel", "BitVoxel.h")
    if os.path.exists(EXE) and os.path.getmtime(EXE) < os.path.getmtime(hdr):
        os.remove(EXE)
    hx = lambda v: bytes(v).hex()  # noqa: E731
    n = len(g["v1"])
    fresh = np.zeros(32, np.uint8)
    lines = []
    for k in range(n):
        lines.append("m %s %s %s %d %d" % (hx(g["v1"][k]), hx(g["v2"][k]), hx(g["coll_in"][k]), g["margin"][k], g["sv_offset"][k]))
        lines.append("m %s %s %s %d %d" % (hx(g["v1"][k]), hx(g["v2"][k]), hx(fresh), g["margin"][k], g["sv_offset"][k]))
        lines.append("s %s %d" % (hx(g["v1"][k]), g["shift"][k]))
    out = subprocess.run([_exe(), "sv"], input="\n".join(lines) + "\n", capture_output=True, text=True, check=True).stdout.split("\n")
    assert len(out) >= 3 * n
    for k in range(n):
        r, c = out[3 * k].split()
        assert int(r) == int(g["result"][k]) and c == hx(g["coll_out"][k]), k
        r, c = out[3 * k + 1].split()
        assert int(r) == int(g["result_fresh"][k]) and c == hx(g["coll_fresh"][k]), k
        assert out[3 * k + 2] == hx(g["shifted"][k]), k


def test_default_collider_thresholds_equal_reference():
    """DefaultCollider::floatToProbability of the facade on the 406 golden thresholds recorded from the reference's
    collider (tests/golden/conversions.npz), the Probability -> float round trip used for the C ABI, and the five
    collide() predicates"""
    g = np.load(os.path.join(ROOT, "tests", "golden", "conversions.npz"))
    for h in ("DefaultCollider.h", "SVCollider.h"):
        hdr = os.path.join(ROOT, "include", "gpu_voxels", "voxel", h)
        if os.path.exists(EXE) and os.path.getmtime(EXE) < os.path.getmtime(hdr):
            os.remove(EXE)
    lines = _run("f2p", g["thr"].view(np.uint32))
    got = np.array([[int(v) for v in ln.split()] for ln in lines if ln not in ("ok", "MISMATCH")])
    want = g["f2p"].astype(np.int32).astype(np.int8).astype(np.int32)   # the reference keeps the low 8 bits
    assert np.array_equal(got[:, 0], want)
    inside = (g["thr"] >= 0) & (g["thr"] <= 1)
    assert got[inside, 1].all()
    assert "MISMATCH" not in lines and "ok" in lines
