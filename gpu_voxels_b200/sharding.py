"""Multi-GPU plumbing of the hot path: one process per GPU, torch.distributed (NCCL over NVLink on the GPUs, gloo in
the CPU tests) for the only exchange the path has -- the reduction of collision results.

The reference is single-GPU. The partition:
  * dense maps: z-slabs. The linear voxel index is z-major (GV/voxelmap/TemplateVoxelMap.h:257-259), so the slab
    [z0,z1) is the contiguous range [z0*dx*dy, z1*dx*dy) and a rank's shard is byte-identical to that slice of the
    unsharded map. Every rank sees the whole cloud and keeps the points of its slab (gvx_map_create_slab).
  * collide map x map / list x map: purely local, then ONE all-reduce of (count | per-meaning counts | 256-bit mask).
  * batch of robot configurations: the environment map is replicated, the BATCH is split; counts are all-gathered.
  * a collide with an offset needs voxels of other ranks' slabs: after connect_map_peers they are read from the
    owners' GPUs through CUDA IPC mappings (collide_with_offset); refused otherwise (GVX_ERR_UNSUPPORTED).
No CUDA code here and no voxel arithmetic: only ranges and collectives.
"""
import numpy as np


def slab_for_rank(dz, rank, world):
    """z-range [z0, z1) of `rank`: dz split as evenly as possible, the first dz % world ranks get one more slice."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world of %d" % (rank, world))
    if world > dz:
        raise ValueError("more ranks (%d) than z-slices (%d)" % (world, dz))
    base, rem = divmod(int(dz), int(world))
    z0 = rank * base + min(rank, rem)
    return z0, z0 + base + (1 if rank < rem else 0)


def slab_bounds(dz, world):
    """world + 1 slab boundaries: rank r owns z in [bounds[r], bounds[r + 1]) (the z_bounds argument of
    gvx_update_logodds_dispatch)"""
    return np.asarray([slab_for_rank(dz, r, world)[0] for r in range(world)] + [int(dz)], np.uint32)


def voxel_range_for_rank(dims, rank, world):
    """linear voxel indices [first, first + n) held by `rank`"""
    z0, z1 = slab_for_rank(dims[2], rank, world)
    plane = int(dims[0]) * int(dims[1])
    return z0 * plane, (z1 - z0) * plane


def batch_range_for_rank(n_cfg, rank, world):
    """configurations [c0, c1) of a planner batch handled by `rank` (may be empty when n_cfg < world)"""
    base, rem = divmod(int(n_cfg), int(world))
    c0 = rank * base + min(rank, rem)
    return c0, c0 + base + (1 if rank < rem else 0)


def list_entries_for_rank(sorted_ids, dims, rank, world, offset=0):
    """[i0, i1) of a sorted voxel-list id array whose map voxel id + offset falls into `rank`'s slab: two binary
    searches, since list ids are sorted (GV/voxellist/TemplateVoxelList.hpp:141-208)."""
    first, n = voxel_range_for_rank(dims, rank, world)
    ids = np.asarray(sorted_ids).astype(np.int64)
    return int(np.searchsorted(ids, first - offset, "left")), int(np.searchsorted(ids, first + n - offset, "left"))


def _dist():
    import torch.distributed as dist
    return dist


def _device_for_backend():
    import torch
    dist = _dist()
    if dist.is_initialized() and dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def allreduce_collision_result(count, mask32=None, per_meaning=None, group=None):
    """Sum the per-rank collision counts (and per-meaning counts) and OR the 256-bit meaning masks over all ranks in
    ONE collective: a single int64 tensor [count | 256 per-meaning counts | 256 mask-bit flags] under SUM (NCCL has
    no bit-OR; bit j of the mask is set iff its summed flag is > 0).
    Returns (count, mask32 or None, per_meaning or None) as python int / numpy arrays."""
    import torch
    dist = _dist()
    buf = torch.zeros(1 + 256 + 256, dtype=torch.int64)
    buf[0] = int(count)
    if per_meaning is not None:
        buf[1:257] = torch.from_numpy(np.asarray(per_meaning, np.uint64).astype(np.int64))
    if mask32 is not None:
        bits = np.unpackbits(np.asarray(mask32, np.uint8).reshape(32), bitorder="little")
        buf[257:513] = torch.from_numpy(bits.astype(np.int64))
    dev = _device_for_backend()
    t = buf.to(dev)
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    out = t.cpu().numpy()
    total = int(out[0])
    pm = out[1:257].astype(np.uint64) if per_meaning is not None else None
    mk = np.packbits((out[257:513] > 0).astype(np.uint8), bitorder="little") if mask32 is not None else None
    return total, mk, pm


def allreduce_count_tensor(count_d, group=None):
    """in-place SUM of a device-resident int64 count (the *_async collide variants write it): no host round trip"""
    dist = _dist()
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(count_d, op=dist.ReduceOp.SUM, group=group)
    return count_d


def allreduce_flag(flag, group=None):
    """OR of the per-rank 'points outside the map' / 'any collision' flags"""
    import torch
    dist = _dist()
    t = torch.tensor([1 if flag else 0], dtype=torch.int32, device=_device_for_backend())
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return bool(t.item())


def allgather_batch_counts(local_counts, n_cfg, group=None):
    """counts of all n_cfg configurations from the per-rank shares of batch_range_for_rank, in configuration order"""
    import torch
    dist = _dist()
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return np.asarray(local_counts, np.uint64)
    width = -(-int(n_cfg) // world)  # equal-sized slots so a plain all_gather works
    dev = _device_for_backend()
    mine = torch.zeros(width, dtype=torch.int64, device=dev)
    lc = np.asarray(local_counts, np.uint64).astype(np.int64)
    mine[: len(lc)] = torch.from_numpy(lc).to(dev)
    parts = [torch.zeros(width, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(parts, mine, group=group)
    out = np.zeros(int(n_cfg), np.uint64)
    for r in range(world):
        c0, c1 = batch_range_for_rank(n_cfg, r, world)
        out[c0:c1] = parts[r].cpu().numpy()[: c1 - c0].astype(np.uint64)
    return out


def _exchange_ipc(create, connect, disconnect, group=None):
    """create() -> (rc, 64-byte handle); handles all-gathered; connect(rank, world, blob) -> rc. True when every rank
    succeeded; on any failure every rank disconnects and the caller keeps the NCCL path."""
    import torch
    dist = _dist()
    if not dist.is_initialized() or dist.get_world_size(group) < 2:
        return False
    dev = _device_for_backend()
    rc, mine = create()
    ok = torch.tensor([1 if rc == 0 else 0], dtype=torch.int32, device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
    if not ok.item():
        return False
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    t = torch.tensor(list(bytes(mine)), dtype=torch.uint8, device=dev)
    parts = [torch.zeros(64, dtype=torch.uint8, device=dev) for _ in range(world)]
    dist.all_gather(parts, t, group=group)
    blob = b"".join(bytes(p.cpu().numpy().tobytes()) for p in parts)
    rc = connect(rank, world, blob)
    ok = torch.tensor([1 if rc == 0 else 0], dtype=torch.int32, device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
    if not ok.item():
        disconnect()
        return False
    dist.barrier(group=group)
    return True


def connect_map_peers(slab_map, group=None):
    """Offset collides on explicit z-slab maps: export this rank's slab of `slab_map` (the map that will be the OTHER side
    of collideWith), all-gather the CUDA IPC handles, map every peer's slab (gvx_map_connect_peers). Afterwards
    `a.collideWith(slab_map, thr, offset)` returns this rank's part of the count -- sum it over the ranks, and keep the
    ranks ordered around the call (collide_with_offset below does both). Returns True when connected."""
    import ctypes as C
    from . import _lib
    dist = _dist()
    if not dist.is_initialized():
        return False
    L = _lib.init()
    world = dist.get_world_size(group)
    bounds = slab_bounds(slab_map.m_dim[2], world)

    def create():
        mine = (C.c_ubyte * 64)()
        return L.gvx_map_ipc_handle(slab_map.handle, mine), mine

    return _exchange_ipc(create,
                         lambda r, w, blob: L.gvx_map_connect_peers(slab_map.handle, r, w, C.c_char_p(blob),
                                                                    bounds.ctypes.data_as(C.c_void_p)),
                         lambda: L.gvx_map_disconnect_peers(slab_map.handle), group)


def collide_with_offset(a, b, coll_threshold, offset, mark=True, group=None):
    """global count of a.collideWith(b, thr, offset) on explicit slab maps whose `b` went through connect_map_peers: every
    rank's earlier work is complete before any rank reads a peer's slab, and every rank's collide is complete before
    anybody continues"""
    import torch
    from . import _lib
    dist = _dist()
    L = _lib.init()
    _lib.check(L.gvx_sync())
    dist.barrier(group=group)
    local = a.collideWith(b, coll_threshold, offset, mark=mark)
    t = torch.tensor([int(local)], dtype=torch.int64, device=_device_for_backend())
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)  # also the barrier behind the reads
    return int(t.item())


def connect_peer_dispatch(capacity_points, group=None):
    """Set up the sensor-frame dispatch (gvx_update_logodds_dispatch / ProbVoxelMap.insertSensorDataDispatch): every rank
    allocates its inboxes (world sources x capacity_points hits, double-buffered), the CUDA IPC handles are exchanged
    with one all_gather, every rank maps its peers. Returns True when connected."""
    import ctypes as C
    from . import _lib
    dist = _dist()
    if not dist.is_initialized():
        return False
    L = _lib.init()
    world = dist.get_world_size(group)

    def create():
        mine = (C.c_ubyte * 64)()
        return L.gvx_peer_dispatch_create(world, int(capacity_points), mine), mine

    return _exchange_ipc(create, lambda r, w, blob: L.gvx_peer_dispatch_connect(r, w, C.c_char_p(blob)),
                         L.gvx_peer_dispatch_disconnect, group)


def connect_peer_mailboxes(group=None):
    """Set up the fused collide + all-reduce path (gvx_collide_*_allreduce): every rank creates its mailbox, the 64-byte
    CUDA IPC handles are exchanged with one all_gather, every rank maps its peers. Returns True when connected; False
    (and the caller keeps using NCCL for the count) when the process group is missing or the IPC mapping is refused."""
    import ctypes as C
    import torch
    from . import _lib
    dist = _dist()
    if not dist.is_initialized() or dist.get_world_size(group) < 2:
        return False
    L = _lib.init()
    mine = (C.c_ubyte * 64)()
    rc = L.gvx_peer_mailbox_create(mine)
    ok = torch.tensor([1 if rc == 0 else 0], dtype=torch.int32, device=_device_for_backend())
    dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
    if not ok.item():
        return False
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = _device_for_backend()
    t = torch.tensor(list(bytes(mine)), dtype=torch.uint8, device=dev)
    parts = [torch.zeros(64, dtype=torch.uint8, device=dev) for _ in range(world)]
    dist.all_gather(parts, t, group=group)
    blob = b"".join(bytes(p.cpu().numpy().tobytes()) for p in parts)
    rc = L.gvx_peer_mailbox_connect(rank, world, C.c_char_p(blob))
    ok = torch.tensor([1 if rc == 0 else 0], dtype=torch.int32, device=dev)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
    if not ok.item():
        L.gvx_peer_mailbox_disconnect()
        return False
    dist.barrier(group=group)
    return True
