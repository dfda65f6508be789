"""Point-cloud file loaders (.xyz / .binvox / .pcd): the product's host-side loader behind the C ABI
(gvx_pointcloud_file_load, no GPU needed) against

  * tests/golden/pointcloud_files.npz -- what the REFERENCE'S OWN readers (GV/helpers/PointcloudFileHandler.cpp,
    BinvoxFileReader.cpp, XyzFileReader.cpp compiled unmodified for the host) return for the committed files under
    tests/golden/files/, malformed ones and the static-variable carry-over between binvox files included;
  * the reference readers run live on every model file of the reference tree (only where /root/reference is mounted);
  * the numpy restatement oracle/gvoracle_files.py -- for .pcd the only check there is (PCL absent: parity unpinned).

Bit-exact: float32 coordinates are compared as raw 32-bit patterns."""
import ctypes as C
import glob
import os
import struct
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FILES = os.path.join(ROOT, "tests", "golden", "files")
GOLD = os.path.join(ROOT, "tests", "golden", "pointcloud_files.npz")
REF_MODELS = "/root/reference/packages/gpu_voxels/models"
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libgvref_files.so")


def same_bits(a, b):
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    return a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32))


def load(path, use_model_path=False, shift=False, offset=None, scaling=1.0):
    from gpu_voxels_b200 import PointcloudFileHandler
    return PointcloudFileHandler.Instance().loadPointCloud(path, use_model_path, None, shift, offset, scaling)


# one fresh process per golden case: the binvox reader's carried-over header values start from zero, as they did in
# the process that recorded the case (tests/golden/make_golden_files.py)
_CHILD = r"""
import sys, numpy as np
sys.path.insert(0, sys.argv[1])
from gpu_voxels_b200 import PointcloudFileHandler
g = np.load(sys.argv[2])
k = int(sys.argv[3])
seq = str(g["seq_%d" % k]).split(",")
shift, scaling, ox, oy, oz = g["args_%d" % k]
h = PointcloudFileHandler.Instance()
for i, name in enumerate(seq):
    if i == len(seq) - 1:
        ok, pts = h.loadPointCloud(name, False, None, bool(shift), (ox, oy, oz), float(scaling))
    else:
        h.loadPointCloud(name, False)
want_ok, want = bool(g["ok_%d" % k]), g["xyz_%d" % k]
if ok != want_ok or (ok and not (pts.shape == want.shape and np.array_equal(pts.view(np.uint32), want.view(np.uint32)))):
    print("MISMATCH", seq, ok, want_ok, pts.shape, want.shape)
    sys.exit(1)
"""


def _n_cases():
    return int(np.load(GOLD)["n_cases"])


@pytest.mark.parametrize("k", range(34))
def test_loader_matches_reference_readers_on_golden_files(k):
    assert _n_cases() == 34
    r = subprocess.run([sys.executable, "-c", _CHILD, ROOT, GOLD, str(k)], cwd=FILES, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr


def test_numpy_restatement_is_pinned_by_the_golden_files():
    from oracle import gvoracle_files as of
    g = np.load(GOLD)
    checked = 0
    for k in range(int(g["n_cases"])):
        seq = str(g["seq_%d" % k]).split(",")
        name = seq[-1]
        if len(seq) != 1 or name not in ("cube.binvox", "grid32.binvox", "values.binvox", "crlf_header.binvox", "simple.xyz"):
            continue
        shift, scaling, ox, oy, oz = g["args_%d" % k]
        data = open(os.path.join(FILES, name), "rb").read()
        pts = of.read_binvox(data) if name.endswith("binvox") else of.read_xyz(data)
        pts = of.post_process(pts, bool(shift), (ox, oy, oz), scaling)
        assert same_bits(pts, g["xyz_%d" % k]), name
        checked += 1
    assert checked >= 10


def test_model_path_prefix_and_unknown_format(monkeypatch, tmp_path):
    monkeypatch.chdir(tmp_path)
    monkeypatch.setenv("GPU_VOXELS_MODEL_PATH", FILES)
    ok, a = load("simple.xyz", use_model_path=True)
    g = np.load(GOLD)
    assert ok and len(a) == 200
    monkeypatch.setenv("GPU_VOXELS_MODEL_PATH", FILES + "/")
    ok2, b = load("simple.xyz", use_model_path=True)
    assert ok2 and same_bits(a, b)
    monkeypatch.delenv("GPU_VOXELS_MODEL_PATH")
    ok3, _ = load("simple.xyz", use_model_path=True)  # variable unset: the bare name is tried (and is not there)
    assert not ok3
    ok4, _ = load(os.path.join(FILES, "cloud.txt"))
    assert not ok4
    from gpu_voxels_b200 import _lib
    assert b"no known file format" in _lib.load().gvx_last_error() or "xyz" in FILES or "pcd" in FILES


@pytest.mark.skipif(not (os.path.isdir(REF_MODELS) and os.path.exists(REF_LIB)), reason="reference tree not mounted")
def test_every_reference_model_file_loads_like_in_the_reference(monkeypatch):
    monkeypatch.chdir(REF_MODELS)
    L = C.CDLL(REF_LIB)
    L.gvrefio_load.restype = C.c_int64
    L.gvrefio_load.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p, C.c_float]
    names = sorted(glob.glob("**/*.binvox", recursive=True) + glob.glob("**/*.xyz", recursive=True))
    assert len(names) > 100
    total = 0
    for i, name in enumerate(names):
        shift, off, scaling = (i % 2 == 1), np.array([0.1 * (i % 3), 0.0, 1.0], np.float32), (1.0 if i % 4 else 0.25)
        # both sides read the SAME sequence of files, so the reader's carried-over header state is the same too
        n = L.gvrefio_load(name.encode(), 0, int(shift), off.ctypes.data, scaling)
        want = np.zeros((max(n, 0), 3), np.float32)
        if n > 0:
            L.gvrefio_fetch(C.c_void_p(want.ctypes.data))
        ok, got = load(name, False, shift, off, scaling)
        assert ok == (n >= 0), name
        assert same_bits(got, want), name
        total += len(got)
    assert total > 1_000_000


# ---- PCD ------------------------------------------------------------------------------------------------------------
def lzf_compress(data: bytes) -> bytes:
    """a small (greedy, hash-less) LZF encoder: literal runs + back references, enough to exercise every token form"""
    out = bytearray()
    lit = bytearray()
    i = 0

    def flush():
        nonlocal lit
        while lit:
            chunk, lit = lit[:32], lit[32:]
            out.append(len(chunk) - 1)
            out.extend(chunk)

    while i < len(data):
        best_len, best_dist = 0, 0
        for dist in (1, 2, 3, 4, 8, 12, 16, 400, 5000):
            if dist > i:
                break
            ln = 0
            while i + ln < len(data) and data[i + ln] == data[i + ln - dist] and ln < 264:
                ln += 1
            if ln > best_len:
                best_len, best_dist = ln, dist
        if best_len >= 3:
            flush()
            ln, d = best_len - 2, best_dist - 1
            if ln < 7:
                out.append((ln << 5) | (d >> 8))
            else:
                out.append((7 << 5) | (d >> 8))
                out.append(ln - 7)
            out.append(d & 0xFF)
            i += best_len
        else:
            lit.append(data[i])
            i += 1
    flush()
    return bytes(out)


def make_pcd(mode, pts, extra=True, comment=True):
    """fields: [normal(F4 x3)] x y [rgb U4] z [curv F8]"""
    n = len(pts)
    rng = np.random.Generator(np.random.PCG64(n))
    if extra:
        dt = np.dtype([("normal", "<f4", (3,)), ("x", "<f4"), ("y", "<f4"), ("rgb", "<u4"), ("z", "<f4"), ("curv", "<f8")])
        head = "FIELDS normal x y rgb z curv\nSIZE 4 4 4 4 4 8\nTYPE F F F U F F\nCOUNT 3 1 1 1 1 1\n"
    else:
        dt = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4")])
        head = "FIELDS x y z\nSIZE 4 4 4\nTYPE F F F\n"  # COUNT is optional
    rec = np.zeros(n, dt)
    rec["x"], rec["y"], rec["z"] = pts[:, 0], pts[:, 1], pts[:, 2]
    if extra:
        rec["normal"] = rng.normal(size=(n, 3))
        rec["rgb"] = rng.integers(0, 2 ** 24, size=n)
        rec["curv"] = rng.uniform(size=n)
    text = ("# .PCD v0.7 - Point Cloud Data file format\n" if comment else "") + "VERSION 0.7\n" + head + \
        "WIDTH %d\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS %d\nDATA %s\n" % (n, n, mode)
    if mode == "ascii":
        def fmt(v):
            return "nan" if v != v else repr(float(v)) if isinstance(v, (float, np.floating)) else str(int(v))
        lines = []
        for r in rec:
            vals = []
            for nm in dt.names:
                vals += [fmt(v) for v in np.atleast_1d(r[nm])]
            lines.append(" ".join(vals))
        return text.encode() + ("\n".join(lines) + "\n").encode()
    if mode == "binary":
        return text.encode() + rec.tobytes()
    soa = b"".join(np.ascontiguousarray(rec[nm]).tobytes() for nm in dt.names)
    packed = lzf_compress(soa)
    return text.encode() + struct.pack("<II", len(packed), len(soa)) + packed


def _pcd_points(n, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    p = rng.uniform(-3, 3, size=(n, 3)).astype(np.float32)
    p[::7] = np.round(p[::7])              # repeated byte patterns -> back references in the LZF stream
    if n > 20:
        p[11] = [np.nan, np.nan, np.nan]   # organised clouds mark invalid pixels like this
    return p


@pytest.mark.parametrize("mode", ["ascii", "binary", "binary_compressed"])
@pytest.mark.parametrize("extra", [False, True])
def test_pcd_reader_against_numpy_restatement(tmp_path, monkeypatch, mode, extra):
    """PARITY UNPINNED (no PCL here): product C++ reader vs the independent numpy reader vs the points written"""
    from oracle import gvoracle_files as of
    monkeypatch.chdir(tmp_path)
    for n in (0, 1, 33, 2000):
        pts = _pcd_points(n, 100 + n)
        data = make_pcd(mode, pts, extra)
        with open("cloud_%d.pcd" % n, "wb") as f:
            f.write(data)
        ok, got = load("cloud_%d.pcd" % n)
        assert ok
        # loadPointCloud always applies scaling * p + offset (1 * p + 0 here), which turns -0.0 into +0.0
        assert same_bits(got, of.post_process(pts))
        assert same_bits(of.read_pcd(data), pts)
        ok, got = load("cloud_%d.pcd" % n, False, True, (0.5, 0.25, 0.125), 0.1)
        if n > 20:   # NaN points: min/shift semantics are the std::min chain of the reference, compare the finite ones
            keep = ~np.isnan(pts[:, 0])
            lo = pts[keep].min(axis=0)
            want = (np.float32(0.1) * (pts - lo)).astype(np.float32) + np.array([0.5, 0.25, 0.125], np.float32)
            assert same_bits(got[keep], want[keep])
        else:
            assert same_bits(got, of.post_process(pts, True, (0.5, 0.25, 0.125), 0.1))


def test_lzf_helper_round_trip():
    from oracle import gvoracle_files as of
    rng = np.random.Generator(np.random.PCG64(5))
    for data in (b"", b"a", b"abcabcabcabcabcabcabc" * 40, bytes(rng.integers(0, 4, size=5000, dtype=np.uint8)), bytes(6000)):
        assert of.lzf_decompress(lzf_compress(data), len(data)) == data


def test_pcd_errors(tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    pts = _pcd_points(10, 1)
    good = make_pcd("binary", pts, True)
    lzf = make_pcd("binary_compressed", pts, True)
    lzf_head = lzf[:lzf.index(b"binary_compressed\n") + len(b"binary_compressed\n")]
    cases = {
        "truncated.pcd": good[:-5],
        "no_data_line.pcd": good[:good.index(b"DATA")],
        "float64_fields.pcd": make_pcd("binary", pts, False).replace(b"SIZE 4 4 4", b"SIZE 8 8 8"),
        "bad_mode.pcd": make_pcd("ascii", pts, False).replace(b"DATA ascii", b"DATA csv"),
        "short_ascii.pcd": make_pcd("ascii", pts, False).rsplit(b"\n", 3)[0] + b"\n",
        # a back reference before any output exists / a stream that expands to fewer bytes than announced
        "bad_lzf.pcd": lzf_head + struct.pack("<II", 2, 10 * 36) + b"\xe0\x00",
        "short_lzf.pcd": lzf_head + struct.pack("<II", 3, 10 * 36) + b"\x01ab",
        "no_z.pcd": make_pcd("binary", pts, False).replace(b"FIELDS x y z", b"FIELDS x y w"),
    }
    for name, data in cases.items():
        with open(name, "wb") as f:
            f.write(data)
        ok, got = load(name)
        assert not ok and len(got) == 0, name
    assert not load("missing.pcd")[0]


def test_abi_rejects_null_arguments():
    from gpu_voxels_b200 import _lib
    L = _lib.load()
    n = C.c_uint64()
    out = C.c_void_p()
    assert L.gvx_pointcloud_file_load(None, 0, 0, None, 1.0, C.byref(out), C.byref(n)) == _lib.ERR_INVALID
    assert L.gvx_pointcloud_file_load(b"a.xyz", 0, 0, None, 1.0, None, C.byref(n)) == _lib.ERR_INVALID
    L.gvx_pointcloud_file_free(None)
    # inserting needs an initialised library: loud failure, not a silent no-op
    assert L.gvx_insert_pointcloud_file(1, os.path.join(FILES, "cloud.txt").encode(), 0, 1, 0, None, 1.0, None) != 0


def test_hostile_headers_cost_nothing(tmp_path, monkeypatch):
    """A header may claim any size: the loaders check it against the bytes that are there before allocating, and no
    exception crosses the C ABI. A .binvox grid of 2 * 10^9 cells with three runs of data yields those runs' voxels."""
    import resource
    import time
    monkeypatch.chdir(tmp_path)
    head = b"# .PCD v0.7\nVERSION 0.7\nFIELDS x y z\nSIZE 4 4 4\nTYPE F F F\nCOUNT 1 1 1\nWIDTH %d\nHEIGHT 1\nPOINTS %d\nDATA %s\n"
    big = 2000000000
    cases = {
        "ascii_claims_2e9.pcd": head % (big, big, b"ascii") + b"1 2 3\n4 5 6\n",
        "binary_claims_2e9.pcd": head % (big, big, b"binary") + b"\0" * 24,
        "lzf_claims_2e9.pcd": head % (big, big, b"binary_compressed") + struct.pack("<II", 4, 0xFFFFFFF0) + b"\x01abc",
        "huge_count.pcd": (head % (2, 2, b"binary")).replace(b"COUNT 1 1 1", b"COUNT 1 1 2000000000") + b"\0" * 24,
    }
    t0 = time.perf_counter()
    rss0 = resource.getrusage(resource.RUSAGE_SELF).ru_maxrss
    for name, data in cases.items():
        with open(name, "wb") as f:
            f.write(data)
        ok, got = load(name)
        assert not ok and len(got) == 0, name
    # binvox: 2000 x 1000 x 1000 cells announced, 3 runs present: 5 voxels, 200 empty cells, 2 voxels
    with open("huge_grid.binvox", "wb") as f:
        f.write(b"#binvox 1\ndim 2000 1000 1000\ntranslate 0 0 0\nscale 1000\ndata\n" + bytes([1, 5, 0, 200, 1, 2]))
    ok, got = load("huge_grid.binvox")
    assert ok and len(got) == 7
    assert np.array_equal(got[:, 1], np.array([0, 1, 2, 3, 4, 205, 206], np.float32)) and not got[:, [0, 2]].any()
    assert time.perf_counter() - t0 < 5.0
    assert resource.getrusage(resource.RUSAGE_SELF).ru_maxrss - rss0 < 200 * 1024  # KiB: nothing of header size was touched
