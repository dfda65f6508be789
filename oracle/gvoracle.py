"""ctypes front-end of the CPU oracle (oracle/gvoracle.c). TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module. The product package gpu_voxels_b200 never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
MAP_BIT, MAP_PROB = 0, 4  # MapType enum values, GV/helpers/common_defines.h:59-73
BITVOXEL_BYTES = 32

_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
_i8p = np.ctypeslib.ndpointer(np.int8, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")


def build(force=False):
    so = os.path.join(_HERE, "libgvoracle.so")
    src = os.path.join(_HERE, "gvoracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "libgvoracle.so"], stdout=subprocess.DEVNULL)
    return so


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        L.gvo_voxel_index_signed.restype = C.c_int64
        L.gvo_create_box_of_points.restype = C.c_uint64
        L.gvo_create_equidistant_points.restype = C.c_uint64
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _dims(d):
    return np.asarray(d, dtype=np.uint32)


def num_threads():
    return lib().gvo_num_threads()


def set_num_threads(n):
    lib().gvo_set_num_threads(int(n))


def map_to_voxels(side, xyz):
    xyz = _f32(xyz).reshape(-1, 3)
    out = np.zeros((len(xyz), 3), np.uint32)
    L = lib()
    for i in range(len(xyz)):
        L.gvo_map_to_voxels(C.c_float(side), _ptr(xyz[i]), _ptr(out[i]))
    return out


def voxel_index_signed(dims, off):
    return lib().gvo_voxel_index_signed(_ptr(_dims(dims)), _ptr(np.asarray(off, np.int32)))


def float_to_probability(thr):
    return lib().gvo_float_to_probability(C.c_float(thr))


def transform_points(M, xyz):
    xyz = _f32(xyz).reshape(-1, 3)
    out = np.empty_like(xyz)
    lib().gvo_transform_points(_ptr(_f32(M).reshape(16)), _ptr(xyz), _ptr(out), C.c_uint64(len(xyz)))
    return out


def new_map(map_type, dims, z0=0, z1=None):
    z1 = dims[2] if z1 is None else z1
    n = int(dims[0]) * int(dims[1]) * int(z1 - z0)
    if map_type == MAP_PROB:
        return np.full(n, -128, np.int8)  # clearMap = memset 0x80, GV/voxelmap/TemplateVoxelMap.hpp:229-246
    return np.zeros(n * BITVOXEL_BYTES, np.uint8)  # GV/voxelmap/TemplateVoxelMap.hpp:205-222


def insert_points(m, map_type, dims, side, xyz, meaning, z0=0, z1=None):
    z1 = dims[2] if z1 is None else z1
    xyz = _f32(xyz).reshape(-1, 3)
    outside = C.c_int(0)
    lib().gvo_insert_points(_ptr(m), map_type, _ptr(_dims(dims)), C.c_uint32(z0), C.c_uint32(z1), C.c_float(side),
                            _ptr(xyz), C.c_uint64(len(xyz)), int(meaning), C.byref(outside))
    return bool(outside.value)


def insert_meta(m, map_type, dims, side, xyz, cloud_sizes, meanings, transforms=None, z0=0, z1=None):
    z1 = dims[2] if z1 is None else z1
    xyz = _f32(xyz).reshape(-1, 3)
    cs = np.asarray(cloud_sizes, np.uint32)
    mg = np.atleast_1d(np.asarray(meanings, np.int32))
    tr = _f32(transforms).reshape(-1) if transforms is not None else None
    outside = C.c_int(0)
    lib().gvo_insert_meta(_ptr(m), map_type, _ptr(_dims(dims)), C.c_uint32(z0), C.c_uint32(z1), C.c_float(side),
                          _ptr(xyz), _ptr(cs), C.c_uint32(len(cs)), _ptr(mg), C.c_uint32(len(mg)), _ptr(tr),
                          C.byref(outside))
    return bool(outside.value)


def update_logodds(m, dims, side, xyz_sensor, pose12, delta=72, z0=0, z1=None, parallel=False):
    """parallel=True: the same update as an OpenMP loop over the points (byte CAS; identical result, see gvoracle.c)"""
    z1 = dims[2] if z1 is None else z1
    xyz = _f32(xyz_sensor).reshape(-1, 3)
    fn = lib().gvo_update_logodds_omp if parallel else lib().gvo_update_logodds
    fn(_ptr(m), _ptr(_dims(dims)), C.c_uint32(z0), C.c_uint32(z1), C.c_float(side), _ptr(xyz),
       C.c_uint64(len(xyz)), _ptr(_f32(pose12).reshape(12)), int(delta))


def insert_sensor_data(m, dims, side, xyz_sensor, pose12, raycast=True, robot_map=None, robot_bit=1, delta_free=-10,
                       delta_occ=72, z0=0, z1=None, counts=False):
    """insertSensorData with ray casting / robot cut in the defined order (all free-space updates, then all hits).
    counts=True: also returns (kf, ko), the per-voxel numbers of decrements and hits."""
    z1 = dims[2] if z1 is None else z1
    xyz = _f32(xyz_sensor).reshape(-1, 3)
    kf = np.zeros(len(m), np.uint32) if counts else None
    ko = np.zeros(len(m), np.uint32) if counts else None
    rc = lib().gvo_insert_sensor_data(_ptr(m), _ptr(_dims(dims)), C.c_uint32(z0), C.c_uint32(z1), C.c_float(side),
                                      _ptr(xyz), C.c_uint64(len(xyz)), _ptr(_f32(pose12).reshape(-1)[:12].copy()),
                                      int(raycast), _ptr(robot_map) if robot_map is not None else None, int(robot_bit),
                                      int(delta_free), int(delta_occ), _ptr(kf) if counts else None,
                                      _ptr(ko) if counts else None)
    assert rc == 0
    return (kf, ko) if counts else None


def collide_map_map(a, a_type, b, b_type, n, thr=1.0, off=0, mark=True):
    cnt = C.c_uint64(0)
    lib().gvo_collide_map_map(_ptr(a), a_type, _ptr(b), b_type, C.c_uint64(n), C.c_float(thr), C.c_int64(off),
                              int(mark), C.byref(cnt))
    return cnt.value


def collide_map_map_thresholds(a, a_type, b, b_type, n, t1, t2, off=0, mark=True):
    """DefaultCollider(Probability t1, Probability t2): int8 thresholds for side a / side b"""
    cnt = C.c_uint64(0)
    lib().gvo_collide_map_map_thresholds(_ptr(a), a_type, _ptr(b), b_type, C.c_uint64(n), int(t1), int(t2),
                                         C.c_int64(off), int(mark), C.byref(cnt))
    return cnt.value


def collide_bit_prob_types(a, b, n, thr=1.0, mark=True, mask=None):
    mask = np.zeros(32, np.uint8) if mask is None else mask
    cnt = C.c_uint64(0)
    lib().gvo_collide_bit_prob_types(_ptr(a), _ptr(b), C.c_uint64(n), C.c_float(thr), int(mark), _ptr(mask),
                                     C.byref(cnt))
    return cnt.value, mask


def clear_bits(a, n, mask):
    lib().gvo_clear_bits(_ptr(a), C.c_uint64(n), _ptr(np.asarray(mask, np.uint8)))


class VoxelList:
    """Host restatement of BitVectorVoxelList storage (GV/voxellist/TemplateVoxelList.h:250-252)."""

    def __init__(self, dims, side):
        self.dims, self.side = _dims(dims), float(side)
        self.ids = np.zeros(0, np.uint32)
        self.coords = np.zeros((0, 3), np.uint32)
        self.voxels = np.zeros((0, 32), np.uint8)

    def __len__(self):
        return len(self.ids)

    def insert_meta(self, xyz, cloud_sizes, meanings):
        xyz = _f32(xyz).reshape(-1, 3)
        cs = np.asarray(cloud_sizes, np.uint32)
        mg = np.atleast_1d(np.asarray(meanings, np.int32))
        cap = len(self.ids) + len(xyz)
        ids = np.zeros(cap, np.uint32); ids[:len(self.ids)] = self.ids
        co = np.zeros((cap, 3), np.uint32); co[:len(self.ids)] = self.coords
        vx = np.zeros((cap, 32), np.uint8); vx[:len(self.ids)] = self.voxels
        size = C.c_uint64(len(self.ids))
        rc = lib().gvo_list_insert_meta(_ptr(ids), _ptr(co), _ptr(vx), C.byref(size), _ptr(self.dims),
                                        C.c_float(self.side), _ptr(xyz), _ptr(cs), C.c_uint32(len(cs)), _ptr(mg),
                                        C.c_uint32(len(mg)))
        assert rc == 0
        n = size.value
        self.ids, self.coords, self.voxels = ids[:n].copy(), co[:n].copy(), vx[:n].copy()

    def insert_points(self, xyz, meaning):
        xyz = _f32(xyz).reshape(-1, 3)
        self.insert_meta(xyz, [len(xyz)], [meaning])

    def collide(self, mode, m, map_type, thr=1.0, off=0, mark=True, mask_in=None, map_first=0, map_n=None):
        n_total = int(self.dims[0]) * int(self.dims[1]) * int(self.dims[2])
        map_n = n_total if map_n is None else map_n
        mask_out = np.zeros(32, np.uint8)
        per = np.zeros(256, np.uint64)
        cnt = C.c_uint64(0)
        mi = np.ascontiguousarray(mask_in, np.uint8) if mask_in is not None else np.zeros(32, np.uint8)
        lib().gvo_collide_list_map(int(mode), _ptr(self.ids), _ptr(self.voxels), C.c_uint64(len(self.ids)), _ptr(m),
                                   map_type, C.c_uint64(n_total), C.c_uint64(map_first), C.c_uint64(map_n),
                                   C.c_float(thr), C.c_int64(off), int(mark), _ptr(mi), _ptr(mask_out), _ptr(per),
                                   C.byref(cnt))
        return cnt.value, mask_out, per


# ---- list x list (numpy restatement) --------------------------------------------------------------------------
def _matching(a, b):
    """findMatchingVoxels, GV/voxellist/BitVoxelList.hpp:408-508: indices (into a, into b) of the entries whose voxel id
    is in both lists (ids are sorted and unique in each list)"""
    _, ia, ib = np.intersect1d(a.ids, b.ids, assume_unique=True, return_indices=True)
    return ia, ib


def list_collide_list(a, b):
    """BitVoxelList::collideWith(BitVectorVoxelList*), BitVoxelList.hpp:86-99: ids of a found in b"""
    return int(np.isin(a.ids, b.ids, assume_unique=True).sum())


def list_collide_list_types(a, b):
    """collideWithTypes(list), BitVoxelList.hpp:101-125 -> (count, OR over matching pairs of (va | vb))"""
    ia, ib = _matching(a, b)
    t = np.zeros(32, np.uint8)
    if len(ia):
        t = np.bitwise_or.reduce(a.voxels[ia] | b.voxels[ib], axis=0)
    return len(ia), t


def list_collide_list_bitcheck(a, b, margin):
    """collideWithBitcheck(list, margin), BitVoxelList.hpp:267-296: BitvectorCollision (margin 0, all bits) or
    BitvectorCollisionWithBitshift(margin, 0) (BitVoxelList.h:40-73) on every matching pair"""
    ia, ib = _matching(a, b)
    if margin == 0:
        return int((a.voxels[ia] & b.voxels[ib]).any(axis=1).sum())
    return sum(bit_margin_check(a.voxels[i], b.voxels[j], margin, 0)[0] for i, j in zip(ia, ib))


def list_collide_list_per_meaning(a, b):
    """collideCountingPerMeaning(list), BitVoxelList.hpp:299-344 -> (total, counts[256]) over a's matching voxels"""
    ia, _ = _matching(a, b)
    bits = np.unpackbits(a.voxels[ia], axis=1, bitorder="little") if len(ia) else np.zeros((0, 256), np.uint8)
    counts = bits.sum(axis=0).astype(np.uint64)
    return int(counts.sum()), counts


def list_subtract(a, b):
    """TemplateVoxelList::subtract, TemplateVoxelList.hpp:609-641: a loses the entries whose id is in b"""
    keep = ~np.isin(a.ids, b.ids, assume_unique=True)
    a.ids, a.coords, a.voxels = a.ids[keep].copy(), a.coords[keep].copy(), a.voxels[keep].copy()


def list_merge(a, b, voxel_offset=(0, 0, 0), new_meaning=None):
    """TemplateVoxelList<BitVoxel>::merge, TemplateVoxelList.hpp:548-600: b's entries appended to a with
    applyOffsetOperator (:68-90: coords + offset and id + getVoxelIndexSigned(offset), both truncated to uint32), their
    voxels replaced by the single bit new_meaning if given; then make_unique (:141-208): stable sort by id, equal ids
    OR-merged, the first entry (a's own, if any) keeps its coordinates"""
    off = np.asarray(voxel_offset, np.int64)
    dx, dy = int(a.dims[0]), int(a.dims[1])
    lin = int(off[2]) * dx * dy + int(off[1]) * dx + int(off[0])
    b_ids = ((b.ids.astype(np.int64) + lin) & 0xFFFFFFFF).astype(np.uint32)
    b_coords = ((b.coords.astype(np.int64) + off) & 0xFFFFFFFF).astype(np.uint32)
    b_vox = b.voxels.copy()
    if new_meaning is not None:
        b_vox[:] = 0
        b_vox[:, new_meaning >> 3] = 1 << (new_meaning & 7)
    ids = np.concatenate([a.ids, b_ids])
    coords = np.concatenate([a.coords, b_coords])
    vox = np.concatenate([a.voxels, b_vox])
    order = np.argsort(ids, kind="stable")
    ids, coords, vox = ids[order], coords[order], vox[order]
    head = np.ones(len(ids), bool)
    head[1:] = ids[1:] != ids[:-1]
    slot = np.cumsum(head) - 1
    out = np.zeros((int(head.sum()), 32), np.uint8)
    np.bitwise_or.at(out, slot, vox)
    a.ids, a.coords, a.voxels = ids[head].copy(), coords[head].copy(), out


def list_shift(a, shift):
    """BitVoxelList::shiftLeftSweptVolumeIDs, BitVoxelList.hpp:560-582"""
    for k in range(len(a.voxels)):
        a.voxels[k] = shift_left(a.voxels[k], shift)


def list_file_bytes(a):
    """TemplateVoxelList::writeToDisk, TemplateVoxelList.hpp:444-478, as bytes"""
    import struct
    return (struct.pack("<i3IfI", 1, int(a.dims[0]), int(a.dims[1]), int(a.dims[2]), np.float32(a.side), len(a.ids))
            + a.ids.astype("<u4").tobytes() + a.coords.astype("<u4").tobytes() + a.voxels.tobytes())


def map_file_bytes(m, map_type, dims, side):
    """TemplateVoxelMap::writeToDisk (bin_mode), GV/voxelmap/TemplateVoxelMap.hpp:666-716, as bytes"""
    import struct
    return struct.pack("<if3I", int(map_type), np.float32(side), int(dims[0]), int(dims[1]), int(dims[2])) + m.tobytes()


def batch_robot_collide(env, env_type, dims, side, base_xyz, cloud_sizes, meanings, transforms, mask=None, thr=1.0):
    base_xyz = _f32(base_xyz).reshape(-1, 3)
    cs = np.asarray(cloud_sizes, np.uint32)
    mg = np.asarray(meanings, np.int32)
    tr = _f32(transforms).reshape(-1, len(cs), 16)
    n_cfg = tr.shape[0]
    n = int(dims[0]) * int(dims[1]) * int(dims[2])
    scratch = np.zeros(n * 32, np.uint8)
    counts = np.zeros(n_cfg, np.uint64)
    mk = np.ascontiguousarray(mask, np.uint8) if mask is not None else None
    rc = lib().gvo_batch_robot_collide(_ptr(env), env_type, _ptr(_dims(dims)), C.c_float(side), _ptr(base_xyz),
                                       _ptr(cs), C.c_uint32(len(cs)), _ptr(mg), _ptr(tr), C.c_uint32(n_cfg),
                                       _ptr(mk), C.c_float(thr), _ptr(scratch), _ptr(counts))
    assert rc == 0
    return counts


def bit_margin_check(v1, v2, margin, sv_offset=0, collisions=None):
    v1, v2 = np.ascontiguousarray(v1, np.uint8), np.ascontiguousarray(v2, np.uint8)
    out = np.zeros(32, np.uint8) if collisions is None else collisions
    r = lib().gvo_bit_margin_check(_ptr(v1), _ptr(v2), _ptr(out), C.c_uint32(margin), C.c_uint32(sv_offset))
    return bool(r), out


def shift_left(v32, shift):
    v = np.ascontiguousarray(v32, np.uint8).copy()
    lib().gvo_shift_left(_ptr(v), C.c_uint32(shift))
    return v


def collide_bitcheck(a, b, n, margin=0, sv_offset=0, mark=True, types=None):
    """collisionCheckBitvector(BitVoxelMap*, SVCollider(1.0, margin), types, sv_offset): (count, types32)"""
    types = np.zeros(32, np.uint8) if types is None else types
    cnt = C.c_uint64(0)
    assert lib().gvo_collide_bitcheck(_ptr(a), _ptr(b), C.c_uint64(n), C.c_uint32(margin), C.c_uint32(sv_offset),
                                      int(mark), _ptr(types), C.byref(cnt)) == 0
    return cnt.value, types


def shift_map(a, n, shift):
    return lib().gvo_shift_map(_ptr(a), C.c_uint64(n), C.c_uint32(shift)) == 0


def dh_chain(dh4, joint_types, joint_values):
    """per-link transforms of a DH chain for a batch of configurations: (n_cfg, n_links, 4, 4) float32"""
    dh4 = _f32(dh4).reshape(-1, 4)
    jt = np.ascontiguousarray(joint_types, np.int32)
    jv = _f32(joint_values).reshape(-1, len(dh4))
    out = np.empty((jv.shape[0], len(dh4), 4, 4), np.float32)
    assert lib().gvo_dh_chain(_ptr(dh4), _ptr(jt), C.c_uint32(len(dh4)), _ptr(jv), C.c_uint32(jv.shape[0]), _ptr(out)) == 0
    return out


def create_box_of_points(mn, mx, delta):
    mn, mx = _f32(mn), _f32(mx)
    n = lib().gvo_create_box_of_points(_ptr(mn), _ptr(mx), C.c_float(delta), None, C.c_uint64(0))
    out = np.zeros((n, 3), np.float32)
    lib().gvo_create_box_of_points(_ptr(mn), _ptr(mx), C.c_float(delta), _ptr(out), C.c_uint64(n))
    return out


def create_equidistant_points(max_points, max_coords, side, checkerboard=False):
    black = np.zeros((max_points, 3), np.float32)
    white = np.zeros((max_points, 3), np.float32) if checkerboard else None
    n = lib().gvo_create_equidistant_points(C.c_uint64(max_points), _ptr(_dims(max_coords)), C.c_float(side),
                                            _ptr(black), _ptr(white))
    return (black[:n], white[:n]) if checkerboard else black[:n]


def bits_of(v32):
    """indices of the set bits of one 256-bit voxel (bit k = byte k>>3, position k&7; GV/helpers/BitVector.h:223-228)"""
    v = np.asarray(v32, np.uint8).reshape(32)
    return [k for k in range(256) if v[k >> 3] & (1 << (k & 7))]
