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
    from _cppbuild import build_cpp, facade_headers
    return build_cpp(SRC, EXE, facade_headers(), opt="-O2", link_gvx=False)


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
    lines = subprocess.run([_exe(), "prob"], capture_output=True, text=True, check=True).stdout.split("\n")
    upd = np.array([[int(v) for v in ln.split()] for ln in lines[:256]], np.int8)
    assert np.array_equal(upd, g["update"])
    assert np.array_equal(np.array([int(v) for v in lines[256].split()], np.int8), g["insert"])
    assert lines[257] == "ok"


def test_bitvector_window_check_and_shift_equal_reference():
    """bitMarginCollisionCheck / performLeftShift of the facade (include/gpu_voxels/voxel/BitVoxel.h) on the 1500 golden
    vectors recorded from the reference's BitVector.h (tests/golden/sweptvolume.npz): results and every output byte"""
    g = np.load(os.path.join(ROOT, "tests", "golden", "sweptvolume.npz"))
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
    lines = _run("f2p", g["thr"].view(np.uint32))
    got = np.array([[int(v) for v in ln.split()] for ln in lines if ln not in ("ok", "MISMATCH")])
    want = g["f2p"].astype(np.int32).astype(np.int8).astype(np.int32)   # the reference keeps the low 8 bits
    assert np.array_equal(got[:, 0], want)
    inside = (g["thr"] >= 0) & (g["thr"] <= 1)
    assert got[inside, 1].all()
    assert "MISMATCH" not in lines and "ok" in lines


# ---- synthetic geometry and rotation matrices (GV/helpers/GeometryGeneration.cpp, cuda_matrices.h) ------------------
def _geom(kind, args):
    r = subprocess.run([_exe(), "geom", kind] + [str(int(a)) for a in args], capture_output=True, check=True)
    return r.stdout


def _clouds(buf, k):
    out, pos = [], 0
    for _ in range(k):
        n = int(np.frombuffer(buf, np.uint64, 1, pos)[0])
        out.append(np.frombuffer(buf, np.float32, 3 * n, pos + 8).reshape(n, 3))
        pos += 8 + 12 * n
    assert pos == len(buf)
    return out


def _split(g, name):
    sizes = g[name + "_sizes"]
    ends = np.cumsum(sizes)
    return [g[name + "_points"][e - s:e] for s, e in zip(sizes, ends)]


def _same_bits(a, b):
    return a.shape == b.shape and np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))


def _geometry_golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "geometry.npz"))


def test_rotation_matrices_equal_reference():
    """Matrix3f::createFromRPY of the facade == the reference's product of single-axis rotations, bit for bit, for 747
    angle triples (special angles, tiny angles, random)"""
    g = _geometry_golden()
    rpy = g["rpy"]
    got = []
    for i in range(0, len(rpy), 150):
        got.append(np.frombuffer(_geom("rpy", rpy[i:i + 150].view(np.uint32).reshape(-1)), np.float32).reshape(-1, 9))
    got = np.concatenate(got)
    assert _same_bits(got, g["rpy_m9"])
    # they are rotations
    m = got.reshape(-1, 3, 3).astype(np.float64)
    assert np.allclose(m @ m.transpose(0, 2, 1), np.eye(3), atol=1e-6)


def test_point_generators_equal_reference():
    """createBoxOfPoints / createSphereOfPoints / createCylinderOfPoints: same points in the same order as the reference's
    translation unit (the loops accumulate `x += delta` in float, so the point SET depends on the rounding)"""
    g = _geometry_golden()
    for kind in ("box", "sphere", "cylinder"):
        want = _split(g, kind)
        assert sum(len(w) for w in want) > 30000
        for p, w in zip(g[kind + "_params"], want):
            got, = _clouds(_geom(kind, p.view(np.uint32)), 1)
            assert _same_bits(got, w), (kind, p)


def test_oriented_boxes_equal_reference():
    """createOrientedBox / createOrientedBoxEdges: the cloud handed to PointCloud::update and the Matrix4f handed to
    transformSelf (roll / pitch / yaw + centre) equal the reference's; the transform itself is the device kernel"""
    g = _geometry_golden()
    want = _split(g, "obox")
    for p, w, m in zip(g["obox_params"], want, g["obox_matrix"]):
        buf = _geom("obox", p.view(np.uint32))
        assert _same_bits(np.frombuffer(buf, np.float32, 16), m), p
        got, = _clouds(buf[64:], 1)
        assert _same_bits(got, w), p
    assert {int(p[10]) for p in g["obox_params"]} == {0, 1}


def test_equidistant_points_and_checkerboards_equal_reference():
    g = _geometry_golden()
    eq, black, white = _split(g, "eq"), _split(g, "cb_black"), _split(g, "cb_white")
    for i, (p, s) in enumerate(zip(g["eq_params"], g["eq_side"])):
        r = subprocess.run([_exe(), "geom", "eq"] + [str(int(v)) for v in p] + [str(int(np.float32(s).view(np.uint32)))],
                           capture_output=True, check=True)
        got = _clouds(r.stdout, 3)
        assert _same_bits(got[0], eq[i]) and _same_bits(got[1], black[i]) and _same_bits(got[2], white[i]), p
    assert len(eq[0]) == 10000  # the reference fixture: 10000 points in the 89 x 123 x 74 map


# ---- value API of Vector3f / Matrix3f / Matrix4f (GV/helpers/cuda_vectors.h:247-331, cuda_matrices.h:83-660) -------
def _mathfile(kind, records, tmp_path):
    path = str(tmp_path / (kind + ".f32"))
    np.ascontiguousarray(records, np.float32).tofile(path)
    return subprocess.run([_exe(), "mathfile", kind, path], capture_output=True, check=True).stdout


def test_rotation_decomposition_equals_reference(tmp_path):
    """Matrix3f::toRPY (solution numbers 0, 1, 2; 330 rotations incl. both pitch singularities) and
    orientationMatrixDiff, bit for bit; and the reference's own property test (GV/test/testing_cudaMath.cu:148-195):
    both decompositions rebuild the rotation to within 0.005"""
    g = _geometry_golden()
    got = np.frombuffer(_mathfile("torpy", g["rot"], tmp_path), np.float32).reshape(-1, 3, 3)
    assert _same_bits(got, g["rot_to_rpy"])
    assert _same_bits(np.frombuffer(_mathfile("diff", g["rot"], tmp_path), np.float32).reshape(-1, 3), g["rot_diff"])
    # property: createFromRPY(toRPY(R)) is R again (both solutions)
    for sol in (0, 1):
        back = np.frombuffer(_geom("rpy", np.ascontiguousarray(got[:150, sol]).view(np.uint32).reshape(-1)), np.float32).reshape(-1, 9)
        assert np.abs(back - g["rot"][:150]).max() < 0.005


def test_matrix_inversion_equals_reference(tmp_path):
    """Gauss-Jordan inversion of Matrix3f / Matrix4f: same result bits, same verdict for singular matrices and for the
    zero-pivot cases that need a row swap; M * inverse(M) is the identity (GV/test/testing_cudaMath.cu:127-146)"""
    g = _geometry_golden()
    for n, kind in ((3, "inv3"), (4, "inv4")):
        got = np.frombuffer(_mathfile(kind, g[kind + "_in"], tmp_path), np.float32).reshape(-1, 1 + n * n)
        assert np.array_equal(got[:, 0] != 0, g[kind + "_ok"] != 0)
        assert _same_bits(got[:, 1:], g[kind + "_out"])
        assert 0 < (g[kind + "_ok"] == 0).sum() < 10
    fixture = g["inv4_in"][90].reshape(4, 4).astype(np.float64)   # the matrix of the reference's test
    inv = g["inv4_out"][90].reshape(4, 4).astype(np.float64)
    assert np.abs(fixture @ inv - np.eye(4)).max() < 2e-6


def test_vector_and_matrix_value_api_equals_reference(tmp_path):
    """length / dot / angleBetween (clamped for parallel vectors) / cross / normalized / abs / normalize / Matrix4f * v /
    compVec / apprx_equal; transpose / a - b / a * b / leftMultiply / getRotation * v / apprx_equal; operator<< texts"""
    g = _geometry_golden()
    rec = np.concatenate([g["vec_a"], g["vec_b"], g["vec_m"]], axis=1)
    got = np.frombuffer(_mathfile("vec", rec, tmp_path), np.float32).reshape(-1, 20)
    assert _same_bits(got, g["vec_ops"])
    assert got[:20, 2].max() < 1e-3 and got[20:40, 2].min() > 3.14 and got[40:50, 19].min() == 1.0 and got[50:60, 19].min() == 1.0 and got[100:, 19].max() == 0.0
    rec = np.concatenate([g["mat_a"], g["mat_b"], g["vec_a"][:90]], axis=1)
    got = np.frombuffer(_mathfile("mat", rec, tmp_path), np.float32).reshape(-1, 53)
    assert _same_bits(got[:, :51], g["mat_ops"])
    assert np.array_equal(got[:, 51] != 0, g["mat_approx"] != 0) and (got[:, 52] == 1.0).all()
    rec = np.concatenate([g["fmt_what"].astype(np.float32)[:, None], g["fmt_in"]], axis=1)
    texts = _mathfile("fmt", rec, tmp_path).split(b"\0")[:-1]
    assert [t.decode() for t in texts] == [str(t) for t in g["fmt_text"]]


def test_reference_bitvector_host_tests_pass_on_the_facade():
    """GV/test/testing_bitvector.cu:36-194 (logic operations, stream operators, setByte, noneButEmpty, left shift of the
    swept-volume IDs) run against include/gpu_voxels/voxel/BitVoxel.h"""
    r = subprocess.run([_exe(), "bitvector"], capture_output=True, text=True, check=True)
    assert r.stdout.strip() == "ok", r.stdout


def test_reference_matrix_host_tests_pass_on_the_facade():
    """GV/test/testing_cudaMath.cu:33-195 (equality, approximate equality, transpose, product, inverse, roll / pitch /
    yaw round trips of 1000 random rotations) run against include/gpu_voxels/helpers/cuda_datatypes.h"""
    r = subprocess.run([_exe(), "cudamath"], capture_output=True, text=True, check=True)
    assert r.stdout.strip() == "ok", r.stdout[-2000:]


def test_option_stand_in_and_boost_aliases():
    """include/icl_core_config/Config.h (the option handling the reference's examples use) and helpers/BoostCompat.h"""
    run = lambda *a: subprocess.run([_exe(), "compat"] + list(a), capture_output=True, text=True, check=True).stdout.strip()
    assert run("--device-identifier=cam7", "--flag") == "cam7|1|1|0"
    assert run("-id", "cam9") == "cam9|1|0|0"
    assert run("--device-identifier", "cam1", "-f") == "cam1|1|1|0"
    assert run() == "|0|0|0"
