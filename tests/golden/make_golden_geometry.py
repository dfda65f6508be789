#!/usr/bin/env python
"""tests/golden/make_golden_geometry.py -- records what the REFERENCE's synthetic-geometry generators and rotation-matrix
constructors return (GV/helpers/GeometryGeneration.cpp, GV/helpers/cuda_matrices.h, compiled for the host into
oracle/_ref/libgvref_geom.so by oracle/Makefile) into tests/golden/geometry.npz. Runs only where /root/reference is
mounted; the .npz travels. tests/test_host_helpers.py compares the facade's generators with it bit for bit.

    make -C oracle ref && python tests/golden/make_golden_geometry.py
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libgvref_geom.so"))
for name in ("box", "sphere", "cylinder", "oriented_box", "equidistant", "checkerboard"):
    getattr(L, "gvrefgeo_" + name).restype = C.c_int64
L.gvrefgeo_box.argtypes = [C.c_void_p, C.c_void_p, C.c_float]
L.gvrefgeo_sphere.argtypes = [C.c_void_p, C.c_float, C.c_float]
L.gvrefgeo_cylinder.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float]
L.gvrefgeo_oriented_box.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_void_p]
L.gvrefgeo_equidistant.argtypes = [C.c_uint64, C.c_void_p, C.c_float]
L.gvrefgeo_checkerboard.argtypes = [C.c_uint64, C.c_void_p, C.c_float, C.c_void_p]
L.gvrefgeo_fetch.argtypes = [C.c_int, C.c_void_p]
L.gvrefgeo_rpy.argtypes = [C.c_float, C.c_float, C.c_float, C.c_void_p]


def f3(v):
    return np.ascontiguousarray(np.asarray(v, np.float32))


def fetch(n, which=0):
    out = np.zeros((n, 3), np.float32)
    if n:
        L.gvrefgeo_fetch(which, out.ctypes.data_as(C.c_void_p))
    return out


def pack(clouds):
    sizes = np.array([len(c) for c in clouds], np.int64)
    return sizes, (np.concatenate(clouds) if clouds else np.zeros((0, 3), np.float32))


def main():
    rng = np.random.Generator(np.random.PCG64(777))
    out = {}

    # rotation matrices: special angles, tiny angles, random angles
    sp = np.array([0.0, -0.0, np.pi, -np.pi, np.pi / 2, -np.pi / 2, np.pi / 4, 1e-8, -3e-5, 2 * np.pi, 7.5, -100.0], np.float32)
    rpy = [np.array([a, b, c], np.float32) for a in sp[:7] for b in sp[:7] for c in (sp[0], sp[4], sp[6])]
    rpy += [rng.choice(sp, 3).astype(np.float32) for _ in range(100)]
    rpy += [rng.uniform(-np.pi, np.pi, 3).astype(np.float32) for _ in range(500)]
    rpy = np.stack(rpy)
    m9 = np.zeros((len(rpy), 9), np.float32)
    for i, a in enumerate(rpy):
        L.gvrefgeo_rpy(float(a[0]), float(a[1]), float(a[2]), m9[i].ctypes.data_as(C.c_void_p))
    out["rpy"], out["rpy_m9"] = rpy, m9

    # boxes of points: the float accumulation of the loop variable is what is pinned
    box_p = np.array([[2.1, 2.1, 2.1, 4.1, 4.1, 4.1, 0.5], [3.1, 3.1, 3.1, 5.1, 5.1, 5.1, 0.5], [1.1, 1.1, 1.1, 3.9, 3.9, 3.9, 0.1],
                      [0, 0, 0, 1, 1, 1, 0.1], [-0.3, 0.2, 0.05, 0.3, 0.25, 0.9, 0.07], [1, 1, 1, 1, 1, 1, 0.5],
                      [2, 2, 2, 1, 3, 3, 0.5], [0.01, 0.02, 0.03, 0.31, 0.32, 0.33, 0.01]], np.float32)
    clouds = []
    for p in box_p:
        n = L.gvrefgeo_box(f3(p[0:3]).ctypes.data_as(C.c_void_p), f3(p[3:6]).ctypes.data_as(C.c_void_p), float(p[6]))
        clouds.append(fetch(n))
    out["box_params"] = box_p
    out["box_sizes"], out["box_points"] = pack(clouds)

    sph_p = np.array([[1, 1, 1, 0.5, 0.1], [0.5, 0.5, 0.5, 0.3, 0.02], [0, 0, 0, 1, 0.25], [2.5, -1.5, 0.75, 0.2, 0.013],
                      [1, 1, 1, 0.0, 0.1], [1, 1, 1, 0.05, 0.1], [0.3, 0.3, 0.3, 0.3, 0.3], [10.24, 5.12, 2.56, 0.64, 0.04]], np.float32)
    clouds = []
    for p in sph_p:
        n = L.gvrefgeo_sphere(f3(p[0:3]).ctypes.data_as(C.c_void_p), float(p[3]), float(p[4]))
        clouds.append(fetch(n))
    out["sphere_params"] = sph_p
    out["sphere_sizes"], out["sphere_points"] = pack(clouds)

    cyl_p = np.array([[1, 1, 1, 0.5, 1.0, 0.1], [0.5, 0.5, 0.5, 0.3, 0.1, 0.02], [0, 0, 0, 1, 3, 0.25], [2.5, -1.5, 0.75, 0.2, 0.5, 0.013],
                      [1, 1, 1, 0.0, 1.0, 0.1], [1, 1, 1, 0.3, 0.0, 0.1], [0.3, 0.3, 0.3, 0.3, 0.7, 0.3]], np.float32)
    clouds = []
    for p in cyl_p:
        n = L.gvrefgeo_cylinder(f3(p[0:3]).ctypes.data_as(C.c_void_p), float(p[3]), float(p[4]), float(p[5]))
        clouds.append(fetch(n))
    out["cylinder_params"] = cyl_p
    out["cylinder_sizes"], out["cylinder_points"] = pack(clouds)

    # oriented boxes: (dim3, center3, rot3, spacing, edges) -> the cloud before the transform and the transform
    ob_p = []
    for edges in (0, 1):
        ob_p += [[0.2, 0.1, 0.3, 1, 2, 3, 0.1, 0.2, 0.3, 0.05, edges], [0.5, 0.5, 0.5, 0, 0, 0, 0, 0, 0, 0.1, edges],
                 [0.13, 0.27, 0.08, -1.5, 0.4, 2.2, -2.0, 1.3, 3.0, 0.031, edges], [0.1, 0.1, 0.1, 5, 5, 5, np.pi / 2, 0, np.pi, 0.1, edges],
                 [0.0, 0.2, 0.2, 1, 1, 1, 0.5, 0.5, 0.5, 0.1, edges]]
    ob_p = np.array(ob_p, np.float32)
    clouds, mats = [], np.zeros((len(ob_p), 16), np.float32)
    for i, p in enumerate(ob_p):
        n = L.gvrefgeo_oriented_box(f3(p[0:3]).ctypes.data_as(C.c_void_p), f3(p[3:6]).ctypes.data_as(C.c_void_p),
                                    f3(p[6:9]).ctypes.data_as(C.c_void_p), float(p[9]), int(p[10]), mats[i].ctypes.data_as(C.c_void_p))
        clouds.append(fetch(n))
    out["obox_params"], out["obox_matrix"] = ob_p, mats
    out["obox_sizes"], out["obox_points"] = pack(clouds)

    # equidistant points / 3-D checkerboards (inputs of the reference's own map tests)
    eq_p = np.array([[10000, 89, 123, 74], [1000, 50, 50, 50], [17, 8, 8, 8], [1, 4, 4, 4], [5000, 128, 64, 32], [64, 4, 4, 4]], np.uint32)
    eq_side = np.array([1.0, 0.5, 1.0, 0.01, 0.02, 1.0], np.float32)
    clouds, black, white = [], [], []
    for p, s in zip(eq_p, eq_side):
        dims = np.ascontiguousarray(p[1:4])
        n = L.gvrefgeo_equidistant(int(p[0]), dims.ctypes.data_as(C.c_void_p), float(s))
        clouds.append(fetch(n))
        nw = C.c_int64(0)
        nb = L.gvrefgeo_checkerboard(int(p[0]), dims.ctypes.data_as(C.c_void_p), float(s), C.byref(nw))
        black.append(fetch(nb, 0))
        white.append(fetch(nw.value, 1))
    out["eq_params"], out["eq_side"] = eq_p, eq_side
    out["eq_sizes"], out["eq_points"] = pack(clouds)
    out["cb_black_sizes"], out["cb_black_points"] = pack(black)
    out["cb_white_sizes"], out["cb_white_points"] = pack(white)

    # ---- value API of Vector3f / Matrix3f / Matrix4f (cuda_vectors.h, cuda_matrices.h) ----
    L.gvrefgeo_to_rpy.argtypes = [C.c_void_p, C.c_uint, C.c_void_p]
    L.gvrefgeo_orientation_diff.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.gvrefgeo_inverse3.argtypes = [C.c_void_p, C.c_void_p]
    L.gvrefgeo_inverse4.argtypes = [C.c_void_p, C.c_void_p]
    L.gvrefgeo_vector_ops.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p]
    L.gvrefgeo_matrix_ops.argtypes = [C.c_void_p, C.c_void_p, C.c_double] + [C.c_void_p] * 6
    L.gvrefgeo_format.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int]
    ptr = lambda a: a.ctypes.data_as(C.c_void_p)
    rot = np.ascontiguousarray(m9[::2][:330])          # rotations incl. the pitch singularities of the special angles
    to_rpy = np.zeros((len(rot), 3, 3), np.float32)     # solution_number 0, 1, 2
    diff = np.zeros((len(rot), 3), np.float32)
    for i in range(len(rot)):
        for sol in range(3):
            L.gvrefgeo_to_rpy(ptr(rot[i]), sol, ptr(to_rpy[i, sol]))
        L.gvrefgeo_orientation_diff(ptr(rot[i]), ptr(rot[(i + 1) % len(rot)]), ptr(diff[i]))
    out["rot"], out["rot_to_rpy"], out["rot_diff"] = rot, to_rpy, diff

    m3s = [rot[k] for k in range(0, 60)] + [rng.normal(0, 2, 9).astype(np.float32) for _ in range(60)]
    m3s += [np.zeros(9, np.float32), np.array([1, 2, 3, 2, 4, 6, 1, 0, 1], np.float32),   # singular
            np.array([0, 1, 0, 1, 0, 0, 0, 0, 1], np.float32), np.array([0, 0, 2, 0, 3, 0, 4, 0, 0], np.float32),  # zero pivots
            np.array([1, 2, 3, 1, 2, 4, 5, 6, 7], np.float32)]                              # zero pivot in the second step
    m3s = np.stack(m3s)
    inv3, ok3 = np.zeros_like(m3s), np.zeros(len(m3s), np.int32)
    for i in range(len(m3s)):
        ok3[i] = L.gvrefgeo_inverse3(ptr(m3s[i]), ptr(inv3[i]))
        if not ok3[i]:
            inv3[i] = 0
    out["inv3_in"], out["inv3_out"], out["inv3_ok"] = m3s, inv3, ok3

    m4s = [mats[k] for k in range(len(mats))] + [rng.normal(0, 2, 16).astype(np.float32) for _ in range(80)]
    m4s += [np.array([0.9751700, -0.218711, -0.0347626, 10, 0.1976770, 0.930432, -0.3085770, 11,
                      0.0998334, 0.294044, 0.9505640, 12, 0, 0, 0, 1], np.float32),            # GV/test/testing_cudaMath.cu fixture
            np.zeros(16, np.float32), np.array([0, 0, 0, 1, 0, 0, 1, 0, 0, 1, 0, 0, 1, 0, 0, 0], np.float32),
            np.array([1, 2, 3, 4, 2, 4, 6, 8, 1, 0, 1, 0, 0, 1, 0, 1], np.float32),
            np.array([1, 2, 3, 4, 1, 2, 4, 5, 0, 0, 0, 7, 3, 1, 2, 2], np.float32)]
    m4s = np.stack(m4s)
    inv4, ok4 = np.zeros_like(m4s), np.zeros(len(m4s), np.int32)
    for i in range(len(m4s)):
        ok4[i] = L.gvrefgeo_inverse4(ptr(m4s[i]), ptr(inv4[i]))
        if not ok4[i]:
            inv4[i] = 0
    out["inv4_in"], out["inv4_out"], out["inv4_ok"] = m4s, inv4, ok4

    va = rng.normal(0, 3, (300, 3)).astype(np.float32)
    vb = rng.normal(0, 3, (300, 3)).astype(np.float32)
    vb[:20] = va[:20] * np.float32(2.5)      # parallel: the acos argument is clamped
    vb[20:40] = -va[20:40]                   # antiparallel
    vb[40:50] = va[40:50]                    # equal
    vb[50:60] = va[50:60] + np.float32(1e-4)  # almost equal (apprx_equal, eps 1e-3)
    vm = np.stack([m4s[i % 90] for i in range(300)])
    vops = np.zeros((300, 20), np.float32)
    for i in range(300):
        L.gvrefgeo_vector_ops(ptr(va[i]), ptr(vb[i]), ptr(vm[i]), 1e-3, ptr(vops[i]))
    out["vec_a"], out["vec_b"], out["vec_m"], out["vec_ops"] = va, vb, vm, vops

    ma, mb = m4s[:90].copy(), np.roll(m4s[:90], 1, axis=0).copy()
    mb[:10] = ma[:10] + np.float32(1e-5)
    mops = np.zeros((90, 16 * 3 + 3), np.float32)
    mapprox = np.zeros(90, np.int32)
    for i in range(90):
        ap = C.c_int(0)
        L.gvrefgeo_matrix_ops(ptr(ma[i]), ptr(mb[i]), 1e-3, ptr(mops[i, 0:16]), ptr(mops[i, 16:32]), ptr(mops[i, 32:48]),
                              ptr(mops[i, 48:51]), ptr(va[i]), C.byref(ap))
        mapprox[i] = ap.value
    out["mat_a"], out["mat_b"], out["mat_ops"], out["mat_approx"] = ma, mb, mops, mapprox

    texts = []
    buf = C.create_string_buffer(4096)
    for what, arr in ((0, va[0]), (0, np.array([1, -2.5, 1e-7], np.float32)), (1, rot[5]), (1, m3s[70]), (2, m4s[3]), (2, m4s[94])):
        L.gvrefgeo_format(what, ptr(np.ascontiguousarray(arr)), buf, 4096)
        texts.append(buf.value.decode())
    out["fmt_what"] = np.array([0, 0, 1, 1, 2, 2], np.int32)
    out["fmt_in"] = np.stack([np.pad(a, (0, 16 - len(a))) for a in (va[0], np.array([1, -2.5, 1e-7], np.float32), rot[5], m3s[70], m4s[3], m4s[94])])
    out["fmt_text"] = np.array(texts)

    path = os.path.join(ROOT, "tests", "golden", "geometry.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: v.shape for k, v in out.items()}, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
