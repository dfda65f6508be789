"""BitVectorVoxelList: host-side mirror of the reference's voxel-list class over the C ABI.

  BitVectorVoxelList   GV/voxellist/BitVoxelList.h:139-150, GV/voxellist/TemplateVoxelList.h

Storage and semantics are the reference's (sorted unique ids, OR-merged 256-bit voxels); all work happens in
libgvx.so (csrc/gvx_list.cu).
"""
import ctypes as C

import numpy as np

from . import _lib
from .api import (BitVectorVoxel, GpuVoxelsMap, MT_BITVECTOR_VOXELLIST, SSIZE_MAX, _log_error, _offset_arg, _points_arg)
from ._lib import MEM_HOST


class BitVectorVoxelList:
    map_type = MT_BITVECTOR_VOXELLIST

    def __init__(self, ref_map_dim, voxel_side_length, map_type=MT_BITVECTOR_VOXELLIST):
        self._L = _lib.init()
        self.m_ref_map_dim = tuple(int(d) for d in ref_map_dim)
        self.m_voxel_side_length = float(voxel_side_length)
        h = C.c_uint64(0)
        _lib.check(self._L.gvx_list_create(C.byref(h), *self.m_ref_map_dim, self.m_voxel_side_length))
        self.handle = h.value

    def destroy(self):
        if getattr(self, "handle", None):
            self._L.gvx_list_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    # -- introspection
    def getMapType(self):
        return self.map_type

    def size(self):
        n = C.c_uint64(0)
        _lib.check(self._L.gvx_list_size(self.handle, C.byref(n)))
        return n.value

    __len__ = size

    def getDimensions(self):
        """the reference reports (number of entries, 1, 1), GV/voxellist/TemplateVoxelList.hpp:418-422"""
        return (self.size(), 1, 1)

    def getVoxelSideLength(self):
        return self.m_voxel_side_length

    def getMemoryUsage(self):
        return self.size() * (4 + 12 + 32)

    def download(self):
        """-> (ids uint32[n], coords uint32[n,3], voxels uint8[n,32])"""
        n = self.size()
        ids, co, vx = np.empty(n, np.uint32), np.empty((n, 3), np.uint32), np.empty((n, 32), np.uint8)
        _lib.check(self._L.gvx_list_download(self.handle, ids.ctypes.data_as(C.c_void_p), co.ctypes.data_as(C.c_void_p),
                                             vx.ctypes.data_as(C.c_void_p)))
        return ids, co, vx

    def upload(self, ids, coords, voxels):
        ids = np.ascontiguousarray(ids, np.uint32)
        vx = np.ascontiguousarray(voxels, np.uint8).reshape(len(ids), 32)
        co = None if coords is None else np.ascontiguousarray(coords, np.uint32).reshape(len(ids), 3)
        _lib.check(self._L.gvx_list_upload(self.handle, ids.ctypes.data_as(C.c_void_p),
                                           co.ctypes.data_as(C.c_void_p) if co is not None else None,
                                           vx.ctypes.data_as(C.c_void_p), len(ids)))

    # -- content
    def clearMap(self):
        _lib.check(self._L.gvx_list_clear(self.handle))

    def insertPointCloud(self, points, voxel_meaning):
        p, n, kind, keep = _points_arg(points)
        _lib.check(self._L.gvx_list_insert_points(self.handle, p, n, kind, int(voxel_meaning)))

    def insertPointCloudFromFile(self, path, use_model_path, voxel_meaning, shift_to_zero=False, offset_XYZ=None,
                                 scaling=1.0):
        """GpuVoxelsMap::insertPointCloudFromFile (GV/GpuVoxelsMap.cpp:36-49) on a list"""
        off, keep = _offset_arg(offset_XYZ)
        rc = self._L.gvx_list_insert_pointcloud_file(self.handle, str(path).encode(), int(bool(use_model_path)),
                                                     int(voxel_meaning), int(bool(shift_to_zero)), off, float(scaling))
        if rc in (_lib.ERR_IO, _lib.ERR_INVALID):
            _log_error(self._L.gvx_last_error().decode())
            return False
        _lib.check(rc)
        return True

    def insertMetaPointCloud(self, meta_point_cloud, voxel_meanings):
        sizes = np.asarray(meta_point_cloud.getPointcloudSizes(), np.uint32)
        mg = np.atleast_1d(np.asarray(voxel_meanings, np.int32))
        xyz = meta_point_cloud.accumulated()
        _lib.check(self._L.gvx_list_insert_meta(self.handle, xyz.ctypes.data_as(C.c_void_p), MEM_HOST,
                                                sizes.ctypes.data_as(C.c_void_p), len(sizes),
                                                mg.ctypes.data_as(C.c_void_p), len(mg)))

    # -- collisions (list x map)
    def _call(self, fn, other, coll_threshold, offset, mark, *extra):
        p, keep = GpuVoxelsMap._off(offset)
        cnt = C.c_uint64(0)
        rc = fn(self.handle, other.handle, coll_threshold, p, int(mark), *extra, C.byref(cnt))
        if rc == _lib.ERR_DIM_MISMATCH:  # GV/voxellist/TemplateVoxelList.hpp:737-742
            _log_error(self._L.gvx_last_error().decode())
            return SSIZE_MAX
        _lib.check(rc)
        return cnt.value

    # -- list x list (GV/voxellist/BitVoxelList.hpp:86-125,267-344; offsets between lists are not supported there)
    def _list_call(self, fn, other, offset, *extra):
        if offset is not None and any(int(v) != 0 for v in offset):
            _log_error("Offset for VoxelList collision not supported! Not performing collision check.")
            return 0
        cnt = C.c_uint64(0)
        rc = fn(self.handle, other.handle, *extra, C.byref(cnt))
        if rc == _lib.ERR_DIM_MISMATCH:
            _log_error(self._L.gvx_last_error().decode())
            return SSIZE_MAX
        _lib.check(rc)
        return cnt.value

    def collideWithBitcheck(self, other, margin=0, offset=None):
        """BitVoxelList::collideWithBitcheck(BitVectorVoxelList*, margin), GV/voxellist/BitVoxelList.hpp:267-296"""
        return self._list_call(self._L.gvx_collide_list_list_bitcheck, other, offset, int(margin))

    def subtract(self, other):
        """TemplateVoxelList::subtract, GV/voxellist/TemplateVoxelList.hpp:609-641"""
        _lib.check(self._L.gvx_list_subtract(self.handle, other.handle))
        return True

    def merge(self, other, voxel_offset=None, new_meaning=None, metric_offset=None):
        """TemplateVoxelList<BitVoxel>::merge (GV/voxellist/TemplateVoxelList.hpp:548-608): append `other`'s entries (ids
        and coordinates shifted by voxel_offset, voxels replaced by the single bit new_meaning if given), make unique.
        metric_offset is converted like mapToVoxelsSigned (floor(c / side) in fp32). Only lists can be merged."""
        if not isinstance(other, BitVectorVoxelList):
            _log_error("THIS OPERATION IS NOT YET SUPPORTED BY THE DATA STRUCTURE!")
            return False
        if metric_offset is not None:
            m = np.asarray(metric_offset, np.float32)
            voxel_offset = np.floor(m / np.float32(self.m_voxel_side_length)).astype(np.int32)
        off = None
        if voxel_offset is not None:
            off = np.ascontiguousarray(voxel_offset, dtype=np.int32).reshape(3)
        _lib.check(self._L.gvx_list_merge(self.handle, other.handle,
                                          off.ctypes.data_as(C.c_void_p) if off is not None else None,
                                          -1 if new_meaning is None else int(new_meaning)))
        return True

    def equals(self, other):
        eq = C.c_int(0)
        _lib.check(self._L.gvx_list_equals(self.handle, other.handle, C.byref(eq)))
        return bool(eq.value)

    def shiftLeftSweptVolumeIDs(self, shift_size):
        """GV/voxellist/BitVoxelList.hpp:560-582: more than 56 bits is logged and nothing happens"""
        rc = self._L.gvx_list_shift_sv_ids(self.handle, int(shift_size))
        if rc == _lib.ERR_INVALID and shift_size > 56:
            _log_error(self._L.gvx_last_error().decode())
            return
        _lib.check(rc)

    def writeToDisk(self, path):
        """GV/voxellist/TemplateVoxelList.hpp:444-478"""
        rc = self._L.gvx_list_write_file(self.handle, str(path).encode())
        if rc != 0:
            _log_error(self._L.gvx_last_error().decode())
        return rc == 0

    def readFromDisk(self, path):
        """GV/voxellist/TemplateVoxelList.hpp:480-526"""
        rc = self._L.gvx_list_read_file(self.handle, str(path).encode())
        if rc != 0:
            _log_error(self._L.gvx_last_error().decode())
        return rc == 0

    def collideWith(self, other, coll_threshold=1.0, offset=None, mark=True):
        if isinstance(other, BitVectorVoxelList):
            return self._list_call(self._L.gvx_collide_list_list, other, offset)
        return self._call(self._L.gvx_collide_list_map, other, coll_threshold, offset, mark)

    def collideWithTypes(self, other, types_in_collision, coll_threshold=1.0, offset=None, mark=True):
        """types_in_collision (BitVectorVoxel) is cleared, then receives the OR of the colliding list voxels"""
        if isinstance(other, BitVectorVoxelList):
            return self._list_call(self._L.gvx_collide_list_list_types, other, offset,
                                   types_in_collision.bytes.ctypes.data_as(C.c_void_p))
        return self._call(self._L.gvx_collide_list_map_types, other, coll_threshold, offset, mark,
                          types_in_collision.bytes.ctypes.data_as(C.c_void_p))

    def collideWithTypeMask(self, other, types_to_check, coll_threshold=1.0, offset=None, mark=True):
        b = np.ascontiguousarray(types_to_check.bytes if isinstance(types_to_check, BitVectorVoxel) else types_to_check,
                                 np.uint8)
        return self._call(self._L.gvx_collide_list_map_mask, other, coll_threshold, offset, mark,
                          b.ctypes.data_as(C.c_void_p))

    def collideCountingPerMeaning(self, other, collisions_per_meaning=None, coll_threshold=1.0, offset=None, mark=True):
        """-> (total, counts[256]); semantics of GV/voxellist/BitVoxelList.hpp:300-344 for list x map"""
        counts = np.zeros(256, np.uint64)
        if isinstance(other, BitVectorVoxelList):
            total = self._list_call(self._L.gvx_collide_list_list_per_meaning, other, offset,
                                    counts.ctypes.data_as(C.c_void_p))
        else:
            total = self._call(self._L.gvx_collide_list_map_per_meaning, other, coll_threshold, offset, mark,
                               counts.ctypes.data_as(C.c_void_p))
        if collisions_per_meaning is not None:
            collisions_per_meaning[:] = counts
        return total, counts
