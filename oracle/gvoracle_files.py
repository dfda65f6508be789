"""oracle/gvoracle_files.py -- TEST INFRASTRUCTURE (never imported by the product path).

numpy restatement of the reference's point-cloud file readers for WELL-FORMED files:

  binvox  BinvoxFileReader::readPointCloud (GV/helpers/BinvoxFileReader.cpp:38-166): header `#binvox 1 / dim d h w /
          translate tx ty tz / scale s / data`, then (value, count) byte pairs; cell order x (depth) outermost, then z
          (height), then y (width); a cell holding exactly 1 becomes the point (s/w * x + tx, s/w * y + ty, s/w * z + tz)
          in fp32, one rounding per operation (:147-160)
  xyz     XyzFileReader::readPointCloud (GV/helpers/XyzFileReader.cpp:28-59): one "x y z" triple per line
  pcd     PcdFileReader::readPointCloud (GV/helpers/PcdFileReader.cpp:31-54) = pcl::io::loadPCDFile<PointXYZ>. PCL is a
          third-party dependency absent from /root/reference and from this image (FIND_PACKAGE(PCL), no version pinned):
          the PCD v0.7 file format is restated from its published description. PARITY UNPINNED for .pcd: there is no
          reference run or reference fixture to check this restatement against; the product's C++ reader and this
          numpy reader are two independent implementations of the same description.
  post    PointcloudFileHandler::loadPointCloud (GV/helpers/PointcloudFileHandler.cpp:124-136): optional shift of the
          minimum corner to zero, then p = scaling * p + offset in fp32

Pinned for binvox / xyz / post by tests/test_pointcloud_files.py against tests/golden/pointcloud_files.npz, which was
produced by the reference's own readers compiled for the host (oracle/ref_files_driver.cpp).
Malformed files (the reference's static-variable and stream-state corner cases) are NOT restated here; those are
covered by the golden fixtures directly."""
import struct

import numpy as np

F32 = np.float32


def read_binvox(data: bytes):
    head, _, payload = data.partition(b"data\n")
    tok = head.split()
    assert tok[0] == b"#binvox"
    dims = tr = scale = None
    i = 2
    while i < len(tok):
        if tok[i] == b"dim":
            dims = [int(t) for t in tok[i + 1:i + 4]]; i += 4
        elif tok[i] == b"translate":
            tr = [F32(float(t)) for t in tok[i + 1:i + 4]]; i += 4
        elif tok[i] == b"scale":
            scale = F32(float(tok[i + 1])); i += 2
        else:
            raise ValueError("restatement covers well-formed headers only")
    depth, height, width = dims
    pairs = np.frombuffer(payload[:len(payload) // 2 * 2], np.uint8).reshape(-1, 2)
    cells = np.repeat(pairs[:, 0], pairs[:, 1].astype(np.int64))
    full = np.zeros(depth * height * width, np.uint8)
    full[:min(len(cells), len(full))] = cells[:len(full)]
    idx = np.nonzero(full == 1)[0]
    x, rem = np.divmod(idx, height * width)
    z, y = np.divmod(rem, width)
    s = F32(scale / F32(width))
    pts = np.empty((len(idx), 3), F32)
    pts[:, 0] = (s * x.astype(F32)).astype(F32) + tr[0]
    pts[:, 1] = (s * y.astype(F32)).astype(F32) + tr[1]
    pts[:, 2] = (s * z.astype(F32)).astype(F32) + tr[2]
    return pts


def read_xyz(data: bytes):
    rows = [[F32(float(t)) for t in line.split()] for line in data.decode().splitlines() if line.strip()]
    return np.array(rows, F32).reshape(-1, 3)


def lzf_decompress(src: bytes, n_out: int) -> bytes:
    out = bytearray()
    i = 0
    while i < len(src):
        ctrl = src[i]; i += 1
        if ctrl < 32:
            out += src[i:i + ctrl + 1]; i += ctrl + 1
        else:
            length = ctrl >> 5
            if length == 7:
                length += src[i]; i += 1
            dist = ((ctrl & 0x1F) << 8) + src[i] + 1; i += 1
            for _ in range(length + 2):
                out.append(out[-dist])
    assert len(out) == n_out
    return bytes(out)


_NP = {("F", 4): "<f4", ("F", 8): "<f8", ("U", 1): "u1", ("U", 2): "<u2", ("U", 4): "<u4", ("I", 1): "i1",
       ("I", 2): "<i2", ("I", 4): "<i4"}


def read_pcd(data: bytes):
    pos = 0
    hdr = {}
    while True:
        eol = data.index(b"\n", pos)
        line = data[pos:eol].decode().strip()
        pos = eol + 1
        if not line or line.startswith("#"):
            continue
        key, *vals = line.split()
        hdr[key] = vals
        if key == "DATA":
            break
    names = hdr["FIELDS"]
    sizes = [int(v) for v in hdr["SIZE"]]
    types = hdr["TYPE"]
    counts = [int(v) for v in hdr.get("COUNT", ["1"] * len(names))]
    n = int(hdr["POINTS"][0]) if "POINTS" in hdr else int(hdr["WIDTH"][0]) * int(hdr.get("HEIGHT", ["1"])[0])
    dt = np.dtype([(nm, _NP[(t, s)], (c,)) for nm, s, t, c in zip(names, sizes, types, counts)])
    mode = hdr["DATA"][0]
    if mode == "ascii":
        rows = [ln.split() for ln in data[pos:].decode().splitlines() if ln.strip()][:n]
        col = np.cumsum([0] + counts)
        k = {nm: col[i] for i, nm in enumerate(names)}
        return np.array([[F32(float(r[k[a]])) for a in "xyz"] for r in rows], F32).reshape(-1, 3)
    if mode == "binary":
        rec = np.frombuffer(data, dt, count=n, offset=pos)
        return np.stack([rec[a][:, 0] for a in "xyz"], axis=1).astype(F32)
    if mode == "binary_compressed":
        packed, plain = struct.unpack_from("<II", data, pos)
        raw = lzf_decompress(data[pos + 8:pos + 8 + packed], plain)
        out, at = {}, 0
        for nm, s, t, c in zip(names, sizes, types, counts):
            out[nm] = np.frombuffer(raw, _NP[(t, s)], count=n * c, offset=at).reshape(n, c)
            at += n * c * s
        return np.stack([out[a][:, 0] for a in "xyz"], axis=1).astype(F32)
    raise ValueError(mode)


def post_process(pts, shift_to_zero=False, offset=(0, 0, 0), scaling=1.0):
    pts = np.array(pts, F32).reshape(-1, 3)
    if shift_to_zero and len(pts):
        pts = pts - pts.min(axis=0)
    return (F32(scaling) * pts).astype(F32) + np.asarray(offset, F32)
