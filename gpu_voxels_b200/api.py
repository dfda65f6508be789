"""Host-side mirror of the reference's public API for the hot path, over the C ABI (libgvx.so).

Same names, argument meaning and error behaviour as the reference's C++ classes so that the parity
tests read like the reference's own Boost tests:

  GpuVoxels            GV/GpuVoxels.h:91-415         (getInstance, initialize, addMap, delMap, getMap, clearMap,
                                                      insertPointCloudIntoMap, insertMetaPointCloudIntoMap)
  BitVectorVoxelMap    GV/voxelmap/BitVoxelMap.h     (collideWith, collideWithTypes, clearBit(s), ...)
  ProbVoxelMap         GV/voxelmap/ProbVoxelMap.h    (collideWith, insertSensorData, ...)
  MetaPointCloud       GV/helpers/MetaPointCloud.h   (host container of concatenated sub-clouds)

Point clouds are numpy float32 (n,3) arrays (host memory) or torch CUDA tensors (device memory, zero copy).
The C++ facade in include/gpu_voxels/ is the drop-in for C++ callers; this module is the same thing for the
Python test and benchmark harness. Nothing here computes voxels on the CPU.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import GvxError, MEM_DEVICE, MEM_HOST

# MapType, GV/helpers/common_defines.h:59-73
MT_BITVECTOR_VOXELMAP, MT_BITVECTOR_VOXELLIST, MT_BITVECTOR_OCTREE, MT_BITVECTOR_MORTON_VOXELLIST = 0, 1, 2, 3
MT_PROBAB_VOXELMAP, MT_PROBAB_VOXELLIST, MT_PROBAB_OCTREE, MT_PROBAB_MORTON_VOXELLIST = 4, 5, 6, 7
MT_COUNTING_VOXELLIST, MT_DISTANCE_VOXELMAP = 8, 9
# BitVoxelMeaning, GV/helpers/common_defines.h:50-58
eBVM_FREE, eBVM_OCCUPIED, eBVM_COLLISION, eBVM_UNKNOWN = 0, 1, 2, 3
eBVM_SWEPT_VOLUME_START, eBVM_SWEPT_VOLUME_END, eBVM_UNDEFINED = 4, 254, 255
BIT_VECTOR_LENGTH = 256
SSIZE_MAX = 2 ** 63 - 1
GPU_VOXELS_MAP_TYPE_NOT_IMPLEMENTED = "THIS TYPE OF DATA STRUCTURE IS NOT YET IMPLEMENTED!"


def _log_error(msg):
    import sys
    print("[gpu_voxels ERROR] " + msg, file=sys.stderr)


def _points_arg(points):
    """-> (pointer, n, mem_kind, keepalive)"""
    if hasattr(points, "data_ptr") and hasattr(points, "is_cuda"):  # torch tensor
        t = points
        if not t.is_cuda:
            t = t.detach().cpu().numpy()
        else:
            import torch
            assert t.dtype == torch.float32 and t.is_contiguous()
            n = t.numel() // 3
            _lib.order_after_torch(t)
            return C.c_void_p(t.data_ptr()), n, MEM_DEVICE, t
        points = t
    a = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 3)
    return a.ctypes.data_as(C.c_void_p), len(a), MEM_HOST, a


def _offset_arg(offset_XYZ):
    if offset_XYZ is None:
        return None, None
    a = np.ascontiguousarray(offset_XYZ, dtype=np.float32).reshape(3)
    return a.ctypes.data_as(C.c_void_p), a


class PointcloudFileHandler:
    """file_handling::PointcloudFileHandler (GV/helpers/PointcloudFileHandler.h:40-100, .cpp:64-137): .xyz / .pcd /
    .binvox files -> points. Host only (no GPU, no gvx_init needed)."""
    _instance = None

    @classmethod
    def Instance(cls):
        if cls._instance is None:
            cls._instance = cls()
        return cls._instance

    def loadPointCloud(self, path, use_model_path, points=None, shift_to_zero=False, offset_XYZ=None, scaling=1.0):
        """-> (ok, float32 array n x 3). `points` (an n x 3 array or None) is what the reference's output vector held
        before the call: the loaded points are appended to it and shift / scaling / offset apply to all of them."""
        if points is not None and len(points):
            raise NotImplementedError("appending to a non-empty vector is only offered by the C++ facade")
        L = _lib.load()
        out, n = C.c_void_p(), C.c_uint64()
        off, keep = _offset_arg(offset_XYZ)
        rc = L.gvx_pointcloud_file_load(str(path).encode(), int(bool(use_model_path)), int(bool(shift_to_zero)), off,
                                        float(scaling), C.byref(out), C.byref(n))
        if rc != 0:
            _log_error(L.gvx_last_error().decode())
            return False, np.zeros((0, 3), np.float32)
        try:
            a = np.ctypeslib.as_array(C.cast(out, C.POINTER(C.c_float)), shape=(n.value, 3)).copy() if n.value else \
                np.zeros((0, 3), np.float32)
        finally:
            L.gvx_pointcloud_file_free(out)
        return True, a


class BitVectorVoxel:
    """256-bit meaning set (GV/voxel/BitVoxel.h:38-108); .bytes is the reference's uint8_t[32] layout."""

    def __init__(self, data=None):
        self.bytes = np.zeros(32, np.uint8) if data is None else np.array(data, np.uint8).reshape(32)

    def setBit(self, k):
        self.bytes[k >> 3] |= np.uint8(1 << (k & 7))

    def clearBit(self, k):
        self.bytes[k >> 3] &= np.uint8(~(1 << (k & 7)) & 0xFF)

    def getBit(self, k):
        return bool(self.bytes[k >> 3] & (1 << (k & 7)))

    def clear(self):
        self.bytes[:] = 0

    def isZero(self):
        return not self.bytes.any()

    def noneButEmpty(self):
        return not ((self.bytes[0] & 0xFE) or self.bytes[1:].any())

    def bits(self):
        return [k for k in range(256) if self.getBit(k)]

    def bitVector(self):
        return self


class MetaPointCloud:
    """Concatenated sub-clouds + sizes (GV/helpers/MetaPointCloud.h:41-241). Host-resident here; the library
    stages it to the device in one copy per insert."""

    def __init__(self, point_clouds):
        self.clouds = [np.ascontiguousarray(c, np.float32).reshape(-1, 3) for c in point_clouds]

    def getNumberOfPointclouds(self):
        return len(self.clouds)

    def getPointcloudSizes(self):
        return [len(c) for c in self.clouds]

    def getAccumulatedPointcloudSize(self):
        return sum(len(c) for c in self.clouds)

    def accumulated(self):
        return np.concatenate(self.clouds) if self.clouds else np.zeros((0, 3), np.float32)

    def syncToDevice(self):
        pass


class GpuVoxelsMap:
    """Base of the dense maps (GV/GpuVoxelsMap.h:50-276, GV/voxelmap/TemplateVoxelMap.h)."""
    map_type = None

    def __init__(self, dim, voxel_side_length, map_type=None, slab=None, _wrap_ptr=None):
        self._L = _lib.init()
        self.m_dim = tuple(int(d) for d in dim)
        self.m_voxel_side_length = float(voxel_side_length)
        h = C.c_uint64(0)
        z0, z1 = (0, self.m_dim[2]) if slab is None else (int(slab[0]), int(slab[1]))
        self.slab = (z0, z1)
        if _wrap_ptr is not None:
            _lib.check(self._L.gvx_map_wrap(C.byref(h), self.map_type, C.c_void_p(_wrap_ptr), *self.m_dim,
                                            self.m_voxel_side_length, z0, z1))
        else:
            _lib.check(self._L.gvx_map_create_slab(C.byref(h), self.map_type, *self.m_dim, self.m_voxel_side_length,
                                                   z0, z1))
        self.handle = h.value

    # -- lifetime
    def destroy(self):
        if getattr(self, "handle", None):
            self._L.gvx_map_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    # -- introspection
    def getMapType(self):
        return self.map_type

    def getDimensions(self):
        return self.m_dim

    def getVoxelSideLength(self):
        return self.m_voxel_side_length

    def getMetricDimensions(self):
        return tuple(np.float32(d) * np.float32(self.m_voxel_side_length) for d in self.m_dim)

    def numVoxels(self):
        return self.m_dim[0] * self.m_dim[1] * (self.slab[1] - self.slab[0])

    def getMemoryUsage(self):
        return self.numVoxels() * (1 if self.map_type == MT_PROBAB_VOXELMAP else 32)

    def getVoidDeviceDataPtr(self):
        p = C.c_void_p()
        _lib.check(self._L.gvx_map_device_ptr(self.handle, C.byref(p), None, None))
        return p.value

    def download(self):
        """raw voxel bytes in the reference layout: int8[n] (Prob) or uint8[n,32] (BitVector)"""
        if self.map_type == MT_PROBAB_VOXELMAP:
            out = np.empty(self.numVoxels(), np.int8)
        else:
            out = np.empty((self.numVoxels(), 32), np.uint8)
        _lib.check(self._L.gvx_map_download(self.handle, out.ctypes.data_as(C.c_void_p)))
        return out

    def upload(self, data):
        a = np.ascontiguousarray(data)
        assert a.nbytes == self.getMemoryUsage()
        _lib.check(self._L.gvx_map_upload(self.handle, a.ctypes.data_as(C.c_void_p)))

    # -- occupancy summary (BitVectorVoxelMap): 1 bit per voxel kept in sync by the library
    def refreshSummary(self):
        _lib.check(self._L.gvx_map_refresh_summary(self.handle))

    def summary(self):
        """-> (bool array of this slab's voxels, valid flag); empty array for maps without a summary"""
        n = self.numVoxels()
        words = np.zeros((n + 31) // 32, np.uint32)
        valid = C.c_int(0)
        _lib.check(self._L.gvx_map_summary_download(self.handle, words.ctypes.data_as(C.c_void_p), len(words),
                                                    C.byref(valid)))
        bits = np.unpackbits(words.view(np.uint8), bitorder="little")[:n].astype(bool)
        return bits, bool(valid.value)

    def coarse(self):
        """-> (bool array, one entry per 4096 voxels of this slab; valid flag): the lazily built coarse mask"""
        n = (self.numVoxels() + 4095) // 4096
        words = np.zeros((n + 31) // 32, np.uint32)
        valid = C.c_int(0)
        _lib.check(self._L.gvx_map_coarse_download(self.handle, words.ctypes.data_as(C.c_void_p), len(words),
                                                   C.byref(valid)))
        return np.unpackbits(words.view(np.uint8), bitorder="little")[:n].astype(bool), bool(valid.value)

    # -- files (GV/voxelmap/TemplateVoxelMap.hpp:666-799): False + a logged error on failure, like the reference
    def writeToDisk(self, path):
        rc = self._L.gvx_map_write_file(self.handle, str(path).encode())
        if rc != 0:
            _log_error(self._L.gvx_last_error().decode())
        return rc == 0

    def readFromDisk(self, path):
        rc = self._L.gvx_map_read_file(self.handle, str(path).encode())
        if rc != 0:
            _log_error(self._L.gvx_last_error().decode())
        return rc == 0

    # -- clearing
    def clearMap(self):
        _lib.check(self._L.gvx_map_clear(self.handle))

    def clearBitVoxelMeaning(self, voxel_meaning):
        rc = self._L.gvx_map_clear_meaning(self.handle, int(voxel_meaning))
        if rc == _lib.ERR_UNSUPPORTED:  # ProbVoxelMap: logged, nothing happens (GV/voxelmap/ProbVoxelMap.hpp:110-116)
            _log_error(self._L.gvx_last_error().decode())
            return
        _lib.check(rc)

    # -- voxelisation
    def insertPointCloud(self, points, voxel_meaning, check_outside=True):
        """returns True if any point was outside the map (the reference only logs a warning)"""
        p, n, kind, keep = _points_arg(points)
        flag = C.c_int(0)
        rc = self._L.gvx_insert_points(self.handle, p, n, kind, int(voxel_meaning),
                                       C.byref(flag) if check_outside else None)
        if rc == _lib.ERR_UNSUPPORTED:  # logged no-op, GV/voxelmap/ProbVoxelMap.hpp:118-140
            _log_error(self._L.gvx_last_error().decode())
            return False
        _lib.check(rc)
        return bool(flag.value)

    def insertPointCloudFromFile(self, path, use_model_path, voxel_meaning, shift_to_zero=False, offset_XYZ=None,
                                 scaling=1.0):
        """GpuVoxelsMap::insertPointCloudFromFile (GV/GpuVoxelsMap.cpp:36-49): False if the file cannot be read"""
        off, keep = _offset_arg(offset_XYZ)
        flag = C.c_int(0)
        rc = self._L.gvx_insert_pointcloud_file(self.handle, str(path).encode(), int(bool(use_model_path)),
                                                int(voxel_meaning), int(bool(shift_to_zero)), off, float(scaling),
                                                C.byref(flag))
        if rc in (_lib.ERR_IO, _lib.ERR_INVALID):
            _log_error(self._L.gvx_last_error().decode())
            return False
        if rc == _lib.ERR_UNSUPPORTED:  # loaded, but the ProbVoxelMap insert is a logged no-op: still true there
            _log_error(self._L.gvx_last_error().decode())
            return True
        _lib.check(rc)
        return True

    def insertMetaPointCloud(self, meta_point_cloud, voxel_meanings, transforms=None, check_outside=True):
        """voxel_meanings: one BitVoxelMeaning or a list with one per sub-cloud (both reference overloads,
        GV/voxelmap/TemplateVoxelMap.hpp:609-664). transforms (optional, n_clouds x 4 x 4): the per-link
        transformSubCloud fused into the insert."""
        mpc = meta_point_cloud
        sizes = np.asarray(mpc.getPointcloudSizes(), np.uint32)
        mg = np.atleast_1d(np.asarray(voxel_meanings, np.int32))
        xyz = mpc.accumulated()
        tr = None if transforms is None else np.ascontiguousarray(transforms, np.float32).reshape(len(sizes), 16)
        flag = C.c_int(0)
        rc = self._L.gvx_insert_meta(self.handle, xyz.ctypes.data_as(C.c_void_p), MEM_HOST,
                                     sizes.ctypes.data_as(C.c_void_p), len(sizes), mg.ctypes.data_as(C.c_void_p),
                                     len(mg), tr.ctypes.data_as(C.c_void_p) if tr is not None else None,
                                     C.byref(flag) if check_outside else None)
        if rc == _lib.ERR_UNSUPPORTED:
            _log_error(self._L.gvx_last_error().decode())
            return False
        _lib.check(rc)
        return bool(flag.value)

    # -- collisions
    @staticmethod
    def _off(offset):
        if offset is None:
            return None, None
        a = np.asarray(offset, np.int32).reshape(3)
        return a.ctypes.data_as(C.c_void_p), a

    def collideWith(self, other, coll_threshold=1.0, offset=None, mark=True):
        """size_t collideWith(const Map*, float coll_threshold = 1.0, const Vector3i& offset = Vector3i())
        (GV/helpers/CollisionInterfaces.h:46,75). mark=True keeps the reference's side effect."""
        p, keep = self._off(offset)
        cnt = C.c_uint64(0)
        _lib.check(self._L.gvx_collide_map_map(self.handle, other.handle, coll_threshold, p, int(mark), C.byref(cnt)))
        return cnt.value

    def collisionCheckWithCounter(self, other, threshold1, threshold2, offset=None, mark=True):
        """collisionCheckWithCounterRelativeTransform(other, DefaultCollider(Probability t1, Probability t2), offset)
        (GV/voxelmap/TemplateVoxelMap.hpp:477-519): two int8 thresholds, one per side"""
        p, keep = self._off(offset)
        cnt = C.c_uint64(0)
        _lib.check(self._L.gvx_collide_map_map_thresholds(self.handle, other.handle, int(threshold1), int(threshold2), p,
                                                          int(mark), C.byref(cnt)))
        return cnt.value

    def collisionCheck(self, other, coll_threshold=1.0):
        flag = C.c_int(0)
        _lib.check(self._L.gvx_collide_map_map_any(self.handle, other.handle, coll_threshold, C.byref(flag)))
        return bool(flag.value)


class BitVectorVoxelMap(GpuVoxelsMap):
    map_type = MT_BITVECTOR_VOXELMAP

    def clearBit(self, bit_index):
        _lib.check(self._L.gvx_map_clear_meaning(self.handle, int(bit_index)))

    def clearBits(self, bits):
        b = np.ascontiguousarray(bits.bytes if isinstance(bits, BitVectorVoxel) else bits, np.uint8)
        _lib.check(self._L.gvx_map_clear_bits(self.handle, b.ctypes.data_as(C.c_void_p)))

    def collisionCheckBitvector(self, other, window_size, types_in_collision, sv_offset=0, mark=True):
        """collisionCheckBitvector(other, SVCollider(1.0, window_size), types, sv_offset), GV/voxelmap/BitVoxelMap.hpp:91-135"""
        cnt = C.c_uint64(0)
        _lib.check(self._L.gvx_collide_map_map_bitcheck(self.handle, other.handle, int(window_size), int(sv_offset),
                                                        int(mark), types_in_collision.bytes.ctypes.data_as(C.c_void_p),
                                                        C.byref(cnt)))
        return cnt.value

    def shiftLeftSweptVolumeIDs(self, shift_size):
        """GV/voxelmap/BitVoxelMap.hpp:234-246"""
        _lib.check(self._L.gvx_map_shift_sv_ids(self.handle, int(shift_size)))

    def collideWithTypes(self, other, types_in_collision, coll_threshold=1.0, offset=None, mark=True):
        """GV/voxelmap/BitVoxelMap.hpp:205-217; types_in_collision (BitVectorVoxel) is OR-ed into; offset ignored
        like in the reference. Against a BitVectorVoxelMap this is the bit-window check with window 0."""
        if isinstance(other, BitVectorVoxelMap):
            return self.collisionCheckBitvector(other, 0, types_in_collision, 0, mark)
        cnt = C.c_uint64(0)
        _lib.check(self._L.gvx_collide_map_map_types(self.handle, other.handle, coll_threshold, int(mark),
                                                     types_in_collision.bytes.ctypes.data_as(C.c_void_p),
                                                     C.byref(cnt)))
        return cnt.value


class ProbVoxelMap(GpuVoxelsMap):
    map_type = MT_PROBAB_VOXELMAP

    def insertSensorData(self, points_sensor_frame, sensor_pose, delta_occupied=72, enable_raycasting=False,
                         cut_real_robot=False, robot_voxel_meaning=eBVM_OCCUPIED, robot_map=None, delta_free=-10):
        """ProbVoxelMap::insertSensorData (GV/voxelmap/ProbVoxelMap.hpp:52-102): the log-odds update of the hit
        voxels, optionally with the free-space ray casting and the robot cut. sensor_pose: 3x4 (or 4x4) row-major
        [R|t]."""
        p, n, kind, keep = _points_arg(points_sensor_frame)
        pose = np.ascontiguousarray(np.asarray(sensor_pose, np.float32).reshape(-1)[:12])
        if not enable_raycasting and not cut_real_robot:
            _lib.check(self._L.gvx_update_logodds(self.handle, p, n, kind, pose.ctypes.data_as(C.c_void_p),
                                                  int(delta_occupied)))
            return
        rh = robot_map.handle if cut_real_robot else 0
        _lib.check(self._L.gvx_insert_sensor_data(self.handle, p, n, kind, pose.ctypes.data_as(C.c_void_p),
                                                  int(enable_raycasting), rh, int(robot_voxel_meaning),
                                                  int(delta_free), int(delta_occupied)))


def _insert_sensor_data_multi(self, frames, sensor_poses, delta_occupied=72):
    """several sensor frames with their own poses in one launch (gvx_update_logodds_multi). frames: list of (n_i,3)
    arrays, or ONE concatenated array / CUDA tensor together with frame_sizes = (sizes list)."""
    if isinstance(frames, tuple):  # (concatenated points, sizes)
        pts, sizes = frames
    else:
        sizes = [len(f) for f in frames]
        pts = np.concatenate([np.ascontiguousarray(f, np.float32).reshape(-1, 3) for f in frames]) if frames else \
            np.zeros((0, 3), np.float32)
    p, n, kind, keep = _points_arg(pts)
    sz = np.asarray(sizes, np.uint32)
    assert int(sz.sum()) == n
    poses = np.ascontiguousarray(np.stack([np.asarray(q, np.float32).reshape(-1)[:12] for q in sensor_poses]))
    _lib.check(self._L.gvx_update_logodds_multi(self.handle, p, kind, sz.ctypes.data_as(C.c_void_p), len(sz),
                                                poses.ctypes.data_as(C.c_void_p), int(delta_occupied)))


ProbVoxelMap.insertSensorDataMulti = _insert_sensor_data_multi


def _insert_sensor_data_dispatch(self, points_sensor_frame, sensor_pose, z_bounds, delta_occupied=72):
    """COLLECTIVE (one process per GPU, z-slab maps, sharding.connect_peer_dispatch done): this rank's sensor frame;
    hits are routed to the rank owning the voxel inside the kernel over NVLink peer memory
    (gvx_update_logodds_dispatch). z_bounds: sharding.slab_bounds(dz, world)."""
    p, n, kind, keep = _points_arg(points_sensor_frame)
    pose = np.ascontiguousarray(np.asarray(sensor_pose, np.float32).reshape(-1)[:12])
    zb = np.ascontiguousarray(z_bounds, np.uint32)
    _lib.check(self._L.gvx_update_logodds_dispatch(self.handle, p, n, kind, pose.ctypes.data_as(C.c_void_p),
                                                   int(delta_occupied), zb.ctypes.data_as(C.c_void_p)))


ProbVoxelMap.insertSensorDataDispatch = _insert_sensor_data_dispatch


GVX_OPT_USE_SUMMARY = 1
GVX_OPT_DH_CHECK_BITS = 2
GVX_OPT_COLLIDE_TMA = 3
GVX_OPT_PDL = 4
GVX_OPT_USE_COARSE = 5
REVOLUTE, PRISMATIC = 0, 1  # robot::DHJointType, GV/robot/dh_robot/KinematicLink.h:37-41


class CollisionLane:
    """A collision lane (include/gvx.h "collision lanes"): read-only collision queries on a stream and result blocks of
    their own, so that queries issued by different host threads overlap on the GPU. The reference's counterpart is its
    per-map locking (GV/voxelmap/TemplateVoxelMap.hpp:329-414, 477-519): threads that collide different maps run
    concurrently. One lane per thread; results equal collideWith(..., mark=False) / collisionCheck."""

    def __init__(self):
        self._L = _lib.init()
        h = C.c_uint64(0)
        _lib.check(self._L.gvx_lane_create(C.byref(h)))
        self.handle = h.value

    def destroy(self):
        if getattr(self, "handle", None):
            self._L.gvx_lane_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    def collideWith(self, a, b, coll_threshold=1.0, offset=None):
        p, keep = GpuVoxelsMap._off(offset)
        cnt = C.c_uint64(0)
        if hasattr(a, "m_ref_map_dim"):  # BitVectorVoxelList x map
            _lib.check(self._L.gvx_lane_collide_list_map(self.handle, a.handle, b.handle, coll_threshold, p, C.byref(cnt)))
        else:
            _lib.check(self._L.gvx_lane_collide_map_map(self.handle, a.handle, b.handle, coll_threshold, p, C.byref(cnt)))
        return cnt.value

    def collisionCheck(self, a, b, coll_threshold=1.0):
        flag = C.c_int(0)
        _lib.check(self._L.gvx_lane_collide_map_map_any(self.handle, a.handle, b.handle, coll_threshold, C.byref(flag)))
        return bool(flag.value)


def set_option(option, value):
    """gvx_set_option: GVX_OPT_USE_SUMMARY = 0 makes every collide stream the voxels (the reference's access pattern)"""
    _lib.check(_lib.init().gvx_set_option(int(option), int(value)))


def transform_points(points, matrix4):
    """PointCloud::transform (kernelTransformCloud, GV/helpers/kernels/MetaPointCloudOperations.cu:77-91)"""
    L = _lib.init()
    a = np.ascontiguousarray(points, np.float32).reshape(-1, 3)
    out = np.empty_like(a)
    M = np.ascontiguousarray(matrix4, np.float32).reshape(16)
    _lib.check(L.gvx_transform_points(a.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), len(a), MEM_HOST,
                                      M.ctypes.data_as(C.c_void_p)))
    return out


def batch_robot_collide(env_map, meta_point_cloud, voxel_meanings, transforms, types_to_check=None, coll_threshold=1.0):
    """counts[c] of the planner's state check (gvl_ompl_planning/gvl_ompl_planner_helper.cpp:169-202) for every
    configuration c in ONE call: robot cleared, link clouds transformed by transforms[c] (n_cfg x n_links x 4 x 4),
    inserted with per-link meanings, collided with env_map. types_to_check: optional BitVectorVoxel collision mask."""
    L = _lib.init()
    sizes = np.asarray(meta_point_cloud.getPointcloudSizes(), np.uint32)
    mg = np.ascontiguousarray(np.broadcast_to(np.asarray(voxel_meanings, np.int32), (len(sizes),)))
    xyz = meta_point_cloud.accumulated()
    tr = np.ascontiguousarray(transforms, np.float32).reshape(-1, len(sizes), 16)
    counts = np.zeros(tr.shape[0], np.uint64)
    mk = None
    if types_to_check is not None:
        mk = np.ascontiguousarray(types_to_check.bytes if isinstance(types_to_check, BitVectorVoxel) else types_to_check,
                                  np.uint8)
    _lib.check(L.gvx_batch_robot_collide(env_map.handle, xyz.ctypes.data_as(C.c_void_p), MEM_HOST,
                                         sizes.ctypes.data_as(C.c_void_p), len(sizes), mg.ctypes.data_as(C.c_void_p),
                                         tr.ctypes.data_as(C.c_void_p), tr.shape[0],
                                         mk.ctypes.data_as(C.c_void_p) if mk is not None else None,
                                         coll_threshold, counts.ctypes.data_as(C.c_void_p)))
    return counts


def _dh_args(dh_params, joint_values):
    """dh_params: per link (d, theta, a, alpha, joint_type) -- robot::DHParameters without the joint value"""
    dh = np.asarray(dh_params, np.float64).reshape(-1, 5)
    dh4 = np.ascontiguousarray(dh[:, :4], np.float32)
    jt = np.ascontiguousarray(dh[:, 4], np.int32)
    jv = np.ascontiguousarray(joint_values, np.float32).reshape(-1, len(dh4))
    return dh4, jt, jv


def dh_chain_transforms(dh_params, joint_values, return_ambiguous=False):
    """KinematicChain::setConfiguration for a batch (GV/robot/dh_robot/KinematicChain.cu:93-126): the matrix each link's
    cloud is transformed by, (n_cfg, n_links, 4, 4) float32, bit-identical to the reference's host arithmetic."""
    L = _lib.init()
    dh4, jt, jv = _dh_args(dh_params, joint_values)
    out = np.empty((jv.shape[0], len(dh4), 4, 4), np.float32)
    amb = C.c_uint32(0)
    _lib.check(L.gvx_dh_chain_transforms(dh4.ctypes.data_as(C.c_void_p), jt.ctypes.data_as(C.c_void_p), len(dh4),
                                         jv.ctypes.data_as(C.c_void_p), jv.shape[0], out.ctypes.data_as(C.c_void_p),
                                         C.byref(amb)))
    return (out, amb.value) if return_ambiguous else out


def batch_robot_collide_dh(env_map, meta_point_cloud, voxel_meanings, dh_params, joint_values, link_cloud=None,
                           types_to_check=None, coll_threshold=1.0):
    """batch_robot_collide with the DH forward kinematics done on the device: joint_values (n_cfg, n_links)"""
    L = _lib.init()
    sizes = np.asarray(meta_point_cloud.getPointcloudSizes(), np.uint32)
    mg = np.ascontiguousarray(np.broadcast_to(np.asarray(voxel_meanings, np.int32), (len(sizes),)))
    xyz = meta_point_cloud.accumulated()
    dh4, jt, jv = _dh_args(dh_params, joint_values)
    lc = np.ascontiguousarray(link_cloud, np.int32) if link_cloud is not None else None
    counts = np.zeros(jv.shape[0], np.uint64)
    mk = None
    if types_to_check is not None:
        mk = np.ascontiguousarray(types_to_check.bytes if isinstance(types_to_check, BitVectorVoxel) else types_to_check,
                                  np.uint8)
    _lib.check(L.gvx_batch_robot_collide_dh(env_map.handle, xyz.ctypes.data_as(C.c_void_p), MEM_HOST,
                                            sizes.ctypes.data_as(C.c_void_p), len(sizes), mg.ctypes.data_as(C.c_void_p),
                                            dh4.ctypes.data_as(C.c_void_p), jt.ctypes.data_as(C.c_void_p), len(dh4),
                                            lc.ctypes.data_as(C.c_void_p) if lc is not None else None,
                                            jv.ctypes.data_as(C.c_void_p), jv.shape[0],
                                            mk.ctypes.data_as(C.c_void_p) if mk is not None else None,
                                            coll_threshold, counts.ctypes.data_as(C.c_void_p)))
    return counts


class GpuVoxels:
    """The singleton facade (GV/GpuVoxels.h:91-415, GV/GpuVoxels.cpp:33-78,164-313,443-559)."""
    _instance = None

    def __init__(self):
        self._L = _lib.init()
        self.m_dim = None
        self.m_voxel_side_length = None
        self.m_managed_maps = {}

    @classmethod
    def getInstance(cls):
        if cls._instance is None:
            cls._instance = GpuVoxels()
        return cls._instance

    def initialize(self, dim_x, dim_y, dim_z, voxel_side_length):
        if self.m_dim is not None:  # second call only warns, GV/GpuVoxels.cpp:52-65
            _log_error("Do not try to initialize GpuVoxels multiple times. Parameters remain unchanged.")
            return
        self.m_dim = (int(dim_x), int(dim_y), int(dim_z))
        self.m_voxel_side_length = float(voxel_side_length)

    def addMap(self, map_type, map_name):
        if map_name in self.m_managed_maps:  # duplicate name -> null shared_ptr, GV/GpuVoxels.cpp:169-176
            _log_error("Map with name '%s' already exists." % map_name)
            return None
        if map_type == MT_BITVECTOR_VOXELMAP:
            m = BitVectorVoxelMap(self.m_dim, self.m_voxel_side_length)
        elif map_type == MT_PROBAB_VOXELMAP:
            m = ProbVoxelMap(self.m_dim, self.m_voxel_side_length)
        elif map_type == MT_BITVECTOR_VOXELLIST:
            from .voxellist import BitVectorVoxelList
            m = BitVectorVoxelList(self.m_dim, self.m_voxel_side_length)
        else:  # the reference throws std::string for unimplemented types, GV/GpuVoxels.cpp:210-211
            raise NotImplementedError(GPU_VOXELS_MAP_TYPE_NOT_IMPLEMENTED)
        self.m_managed_maps[map_name] = m
        return m

    def delMap(self, map_name):
        m = self.m_managed_maps.pop(map_name, None)
        if m is None:
            _log_error("Map with name '%s' not found." % map_name)
            return False
        m.destroy()
        return True

    def getMap(self, map_name):
        m = self.m_managed_maps.get(map_name)
        if m is None:
            _log_error("Map with name '%s' not found." % map_name)
        return m

    def clearMap(self, map_name, voxel_meaning=None):
        m = self.getMap(map_name)
        if m is None:
            return False
        if voxel_meaning is None:
            m.clearMap()
        else:
            m.clearBitVoxelMeaning(voxel_meaning)
        return True

    def insertPointCloudIntoMap(self, cloud, map_name, voxel_meaning):
        m = self.getMap(map_name)
        if m is None:
            return False
        m.insertPointCloud(cloud, voxel_meaning)
        return True

    def insertPointCloudFromFile(self, map_name, path, use_model_path, voxel_meaning, shift_to_zero=False,
                                 offset_XYZ=None, scaling=1.0):
        """GpuVoxels::insertPointCloudFromFile (GV/GpuVoxels.cpp:427-440)"""
        m = self.getMap(map_name)
        if m is None:
            return False
        return m.insertPointCloudFromFile(path, use_model_path, voxel_meaning, shift_to_zero, offset_XYZ, scaling)

    def insertMetaPointCloudIntoMap(self, meta_point_cloud, map_name, voxel_meanings):
        m = self.getMap(map_name)
        if m is None:
            return False
        m.insertMetaPointCloud(meta_point_cloud, voxel_meanings)
        return True

    def getDimensions(self):
        return self.m_dim

    def getVoxelSideLength(self):
        return self.m_voxel_side_length
