#!/usr/bin/env python
"""tests/golden/make_golden_files.py -- point-cloud FILE fixtures and what the REFERENCE'S OWN READERS return for them.

Writes small synthetic .binvox / .xyz files (well-formed and malformed) into tests/golden/files/ and records, in
tests/golden/pointcloud_files.npz, the result of the reference's PointcloudFileHandler::loadPointCloud for each case:
oracle/_ref/libgvref_files.so is GV/helpers/{PointcloudFileHandler,BinvoxFileReader,XyzFileReader}.cpp compiled
unmodified for the host (oracle/ref_files_driver.cpp, oracle/Makefile). Run in the authoring container:

    make -C oracle && python tests/golden/make_golden_files.py

A "case" is a SEQUENCE of loads in one fresh process state (the reference's binvox reader keeps header values in
function-local statics, so what a file without `translate` / `scale` yields depends on the file read before it); the
result of the LAST load of the sequence is recorded. tests/test_pointcloud_files.py replays the sequences through the
product's C ABI (gvx_pointcloud_file_load) and compares bit for bit; it needs neither /root/reference nor a GPU.
PCD files are not covered here: the reference reads them through PCL, which is not available (see that test).
Deterministic (seeded)."""
import ctypes as C
import os
import struct
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
FILES = os.path.join(HERE, "files")


def rle(cells):
    """binvox run-length pairs (value, count <= 255) of a flat uint8 array"""
    out = bytearray()
    i = 0
    while i < len(cells):
        j = i
        while j < len(cells) and cells[j] == cells[i] and j - i < 255:
            j += 1
        out += bytes((int(cells[i]), j - i))
        i = j
    return bytes(out)


def binvox(header_lines, payload):
    return ("\n".join(header_lines) + "\ndata\n").encode() + payload


def write_files():
    os.makedirs(FILES, exist_ok=True)
    rng = np.random.Generator(np.random.PCG64(4711))
    f = {}
    cube = (rng.uniform(size=6 * 5 * 7) < 0.3).astype(np.uint8)
    f["cube.binvox"] = binvox(["#binvox 1", "dim 6 5 7", "translate -0.527372 0.25 1e-3", "scale 2.54884"], rle(cube))
    big = (rng.uniform(size=32 * 32 * 32) < 0.05).astype(np.uint8)
    f["grid32.binvox"] = binvox(["#binvox 1", "dim 32 32 32", "translate 0.1 -7.25 3", "scale 0.77"], rle(big))
    vals = rng.integers(0, 4, size=4 * 4 * 4).astype(np.uint8)  # run values 2 and 3 are not voxels
    f["values.binvox"] = binvox(["#binvox 1", "dim 4 4 4", "translate 0 0 0", "scale 1"], rle(vals))
    # header without translate / with scale BEFORE dim: inherits from the file read before
    f["no_translate.binvox"] = binvox(["#binvox 1", "dim 3 3 3", "scale 3"], rle(np.ones(27, np.uint8)))
    f["scale_first.binvox"] = binvox(["#binvox 1", "scale 12", "dim 2 3 4", "translate 1 2 3"], rle(np.ones(24, np.uint8)))
    f["no_scale.binvox"] = binvox(["#binvox 1", "dim 2 2 2", "translate 0.5 0.5 0.5"], rle(np.ones(8, np.uint8)))
    f["unknown_keyword.binvox"] = binvox(["#binvox 1", "author somebody with spaces", "dim 2 2 2", "translate 0 0 0", "scale 2"],
                                         rle(np.array([1, 0, 0, 1, 1, 1, 0, 1], np.uint8)))
    f["short_data.binvox"] = binvox(["#binvox 1", "dim 4 4 4", "translate 0 0 0", "scale 4"], rle(np.ones(20, np.uint8)))
    f["odd_tail.binvox"] = binvox(["#binvox 1", "dim 4 4 4", "translate 0 0 0", "scale 4"], rle(np.ones(20, np.uint8)) + b"\x01")
    f["overflow.binvox"] = binvox(["#binvox 1", "dim 2 2 2", "translate 0 0 0", "scale 2"], bytes((1, 9)))
    f["crlf_header.binvox"] = ("#binvox 1\r\ndim 2 2 2\r\ntranslate 1 1 1\r\nscale 2\r\ndata\n").encode() + rle(np.ones(8, np.uint8))
    f["bad_magic.binvox"] = b"#binvax 1\ndim 2 2 2\ndata\n" + bytes((1, 8))
    f["no_dim.binvox"] = b"#binvox 1\ntranslate 0 0 0\nscale 1\ndata\n" + bytes((1, 8))
    f["no_data.binvox"] = b"#binvox 1\ndim 2 2 2\ntranslate 0 0 0\nscale 1\n"
    f["bad_dim.binvox"] = b"#binvox 1\ndim 2 two 2\ntranslate 0 0 0\nscale 1\ndata\n" + bytes((1, 8))
    f["empty.binvox"] = b""
    f["zero_dim.binvox"] = binvox(["#binvox 1", "dim 0 4 4", "translate 0 0 0", "scale 1"], b"")
    pts = rng.uniform(-2, 2, size=(200, 3))
    f["simple.xyz"] = "".join("%.9g %.9g %.9g\n" % tuple(p) for p in pts).encode()
    f["messy.xyz"] = (
        "1 2 3\n"
        "  4.5\t-5e-1   6E2  7 8 9   10 11\n"      # two triples and an incomplete one
        "\n"
        "12 13 14\r\n"                               # CR before LF is white space
        "15 16 abc 17 18 19\n"                       # line abandoned at 'abc'
        "nan 1 2\n"                                  # not a number for the stream extractor
        "inf 1 2\n"
        "0x10 1 2 3 4 5\n"                           # '0' then 'x10' fails
        "1,2,3\n"
        "1e50 2 3\n"                                 # overflows float: extraction fails
        "1e-50 2 3\n"                                # underflows
        ".5 -.25 +3.\n"
        "1.0f 2 3\n"
        "20 21 22 # comment 23 24 25\n"
        "-0 0 -0.0\n"
        "3.4028235e38 -3.4028235e38 1.17549435e-38\n"
        "1e-45 7e-46 1e-46\n"
        "0.1 0.2 0.3").encode()                      # no newline at the end
    f["empty.xyz"] = b""
    f["cloud.txt"] = b"1 2 3\n"                     # no known format
    f["xyz_named.binvox"] = f["cube.binvox"]         # "xyz" in the name wins: read as text
    f["grid_xyz.pcd"] = b"1 2 3\n4 5 6\n"           # likewise: never reaches the PCD reader
    for name, data in f.items():
        with open(os.path.join(FILES, name), "wb") as fh:
            fh.write(data)
    return sorted(f)


# (sequence of files, shift_to_zero, offset, scaling); the last file's result is recorded
CASES = [
    (["cube.binvox"], 0, (0, 0, 0), 1.0),
    (["cube.binvox"], 1, (0, 0, 0), 1.0),
    (["cube.binvox"], 1, (0.1, -0.2, 0.3), 0.01),
    (["cube.binvox"], 0, (1.5, 2.5, -3.5), 3.3),
    (["grid32.binvox"], 0, (0, 0, 0), 1.0),
    (["grid32.binvox"], 1, (0.005, 0.005, 0.005), 0.04),
    (["values.binvox"], 0, (0, 0, 0), 1.0),
    (["no_translate.binvox"], 0, (0, 0, 0), 1.0),                       # fresh process: translate 0
    (["cube.binvox", "no_translate.binvox"], 0, (0, 0, 0), 1.0),        # inherits cube's translate
    (["scale_first.binvox"], 0, (0, 0, 0), 1.0),                        # fresh: scale / 0
    (["cube.binvox", "scale_first.binvox"], 0, (0, 0, 0), 1.0),         # scale / cube's width
    (["grid32.binvox", "no_scale.binvox"], 0, (0, 0, 0), 1.0),
    (["unknown_keyword.binvox"], 0, (0, 0, 0), 1.0),
    (["short_data.binvox"], 0, (0, 0, 0), 1.0),
    (["odd_tail.binvox"], 0, (0, 0, 0), 1.0),
    (["overflow.binvox"], 0, (0, 0, 0), 1.0),
    (["crlf_header.binvox"], 0, (0, 0, 0), 1.0),
    (["bad_magic.binvox"], 0, (0, 0, 0), 1.0),
    (["no_dim.binvox"], 0, (0, 0, 0), 1.0),
    (["no_data.binvox"], 0, (0, 0, 0), 1.0),
    (["bad_dim.binvox"], 0, (0, 0, 0), 1.0),
    (["empty.binvox"], 0, (0, 0, 0), 1.0),
    (["zero_dim.binvox"], 0, (0, 0, 0), 1.0),
    (["simple.xyz"], 0, (0, 0, 0), 1.0),
    (["simple.xyz"], 1, (0, 0, 0), 1.0),
    (["simple.xyz"], 1, (2.56, 2.56, 0.0), 0.5),
    (["messy.xyz"], 0, (0, 0, 0), 1.0),
    (["messy.xyz"], 1, (1, 1, 1), 2.0),
    (["empty.xyz"], 0, (0, 0, 0), 1.0),
    (["cloud.txt"], 0, (0, 0, 0), 1.0),
    (["xyz_named.binvox"], 0, (0, 0, 0), 1.0),
    (["grid_xyz.pcd"], 0, (0, 0, 0), 1.0),
    (["does_not_exist.xyz"], 0, (0, 0, 0), 1.0),
    (["does_not_exist.binvox"], 0, (0, 0, 0), 1.0),
]

_CHILD = r"""
import ctypes as C, sys, numpy as np
lib, out = sys.argv[1], sys.argv[2]
seq, shift, scaling = sys.argv[3].split(","), int(sys.argv[4]), float(sys.argv[5])
off = np.array([float(v) for v in sys.argv[6].split(",")], np.float32)
L = C.CDLL(lib)
L.gvrefio_load.restype = C.c_int64
L.gvrefio_load.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_float]
n = -1
for i, name in enumerate(seq):
    last = i == len(seq) - 1
    n = L.gvrefio_load(name.encode(), 0, shift if last else 0, off.ctypes.data if last else np.zeros(3, np.float32).ctypes.data,
                       scaling if last else 1.0)
a = np.zeros((max(n, 0), 3), np.float32)
if n > 0:
    L.gvrefio_fetch(C.c_void_p(a.ctypes.data))
np.save(out, a)
sys.exit(0 if n >= 0 else 3)
"""


def main():
    import subprocess
    import tempfile
    lib = os.path.join(ROOT, "oracle", "_ref", "libgvref_files.so")
    names = write_files()
    out = {"n_cases": np.array(len(CASES))}
    with tempfile.TemporaryDirectory() as tmp:
        for k, (seq, shift, off, scaling) in enumerate(CASES):
            res = os.path.join(tmp, "r.npy")
            # a fresh process per case: the reference's statics start from zero; cwd = the files directory, so that
            # the path the readers see is the bare file name (the format is detected by substring of the WHOLE path)
            r = subprocess.run([sys.executable, "-c", _CHILD, lib, res, ",".join(seq), str(shift), repr(scaling),
                                ",".join(repr(float(v)) for v in off)], cwd=FILES, stderr=subprocess.DEVNULL)
            assert r.returncode in (0, 3), (seq, r.returncode)
            out["ok_%d" % k] = np.array(r.returncode == 0)
            out["xyz_%d" % k] = np.load(res)
            out["seq_%d" % k] = np.array(",".join(seq))
            out["args_%d" % k] = np.array([shift, scaling, off[0], off[1], off[2]], np.float64)
            print("%-45s ok=%d n=%d" % ("+".join(seq), r.returncode == 0, len(out["xyz_%d" % k])))
    np.savez_compressed(os.path.join(HERE, "pointcloud_files.npz"), **out)
    print("wrote", len(names), "files and", len(CASES), "cases")


def math_helpers():
    """computeLinearLoad / interpolateLinear of the reference (GV/helpers/MathHelpers.cpp compiled for the host into
    oracle/_ref/libgvref_files.so) -> tests/golden/math_helpers.npz"""
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libgvref_files.so"))
    L.gvrefio_lerp_f.restype, L.gvrefio_lerp_f.argtypes = C.c_float, [C.c_float] * 3
    L.gvrefio_lerp_d.restype, L.gvrefio_lerp_d.argtypes = C.c_double, [C.c_double] * 3
    rng = np.random.Generator(np.random.PCG64(99))
    n = np.concatenate([[0, 1, 2, 1023, 1024, 1025, 2047, 2048, 65535 * 1024 - 1, 65535 * 1024, 65535 * 1024 + 1, 2 ** 31, 2 ** 32 - 1],
                        rng.integers(0, 2 ** 32, size=200)]).astype(np.uint32)
    load = np.zeros((len(n), 2), np.uint32)
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(2)
    os.dup2(devnull, 2)  # the n == 0 case logs an error
    for i, v in enumerate(n):
        b, t = C.c_uint32(), C.c_uint32()
        L.gvrefio_linear_load(C.c_uint32(int(v)), C.byref(b), C.byref(t))
        load[i] = (b.value, t.value)
    os.dup2(saved, 2)
    f = (rng.normal(size=(400, 3)) * 10.0 ** rng.integers(-3, 4, size=(400, 1))).astype(np.float32)
    f[:, 2] = rng.uniform(-0.5, 1.5, size=400).astype(np.float32)
    f[0] = (1, 2, 0); f[1] = (1, 2, 1); f[2] = (0.1, 0.2, 0.5); f[3] = (3.4e38, 3.4e38, 0.5); f[4] = (1e-45, 1e-45, 0.5)
    lf = np.array([L.gvrefio_lerp_f(*[float(x) for x in row]) for row in f], np.float32)
    d = f.astype(np.float64) * rng.uniform(0.9, 1.1, size=(400, 3))
    ld = np.array([L.gvrefio_lerp_d(*[float(x) for x in row]) for row in d], np.float64)
    np.savez_compressed(os.path.join(HERE, "math_helpers.npz"), n=n, load=load, f=f, lerp_f=lf, d=d, lerp_d=ld)
    print("math helpers:", len(n), "launch shapes,", len(f), "+", len(d), "interpolations")


def prob_voxel():
    """ProbabilisticVoxel::updateOccupancy for all 256 x 256 (occupancy, delta) pairs and ::insert for all 256 meanings,
    from the reference's own header (GV/voxel/ProbabilisticVoxel.hpp via oracle/_ref/libgvref_cpu.so)"""
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libgvref_cpu.so"))
    upd = np.array([[L.gvrefcpu_update_occupancy(o, d) for d in range(-128, 128)] for o in range(-128, 128)], np.int8)
    ins = np.zeros(256, np.int8)
    for m in range(256):
        cell = np.array([5], np.int8)
        L.gvrefcpu_insert_at(C.c_void_p(cell.ctypes.data), 1, C.c_uint64(0), m)
        ins[m] = cell[0]
    np.savez_compressed(os.path.join(HERE, "prob_voxel.npz"), update=upd, insert=ins)
    print("prob voxel: 65536 updates,", int((ins == 127).sum()), "meanings -> 127,", int((ins == -127).sum()), "-> -127")


if __name__ == "__main__":
    main()
    math_helpers()
    prob_voxel()
