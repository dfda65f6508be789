"""ctypes binding of libgvx.so (the C ABI declared in include/gvx.h).

The product path: there is NO CPU fallback. If the CUDA library is missing or no GPU is present,
loading / gvx_init fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libgvx.so")

GVX_OK = 0
ERR_INVALID, ERR_CUDA, ERR_NOMEM, ERR_UNSUPPORTED, ERR_DIM_MISMATCH, ERR_NOT_INITIALIZED, ERR_IO = 1, 2, 3, 4, 5, 6, 7
MEM_HOST, MEM_DEVICE, MEM_HOST_PINNED = 0, 1, 2

u64, u32, i32, f32, vp = C.c_uint64, C.c_uint32, C.c_int32, C.c_float, C.c_void_p
P = C.POINTER

# name -> (restype, argtypes); must list every symbol include/gvx.h declares (checked by tests/test_abi.py)
SIGNATURES = {
    "gvx_init": (C.c_int, [C.c_int]),
    "gvx_shutdown": (C.c_int, []),
    "gvx_last_error": (C.c_char_p, []),
    "gvx_set_stream": (C.c_int, [vp]),
    "gvx_sync": (C.c_int, []),
    "gvx_wait_event": (C.c_int, [vp]),
    "gvx_device_info": (C.c_int, [P(C.c_int), P(u64), P(u64), C.c_char_p]),
    "gvx_launch_count": (u64, []),
    "gvx_set_l2_fetch_granularity": (C.c_int, [C.c_int, P(C.c_int)]),
    "gvx_buffer_create": (C.c_int, [P(u64), u64]),
    "gvx_buffer_destroy": (C.c_int, [u64]),
    "gvx_buffer_upload": (C.c_int, [u64, u64, vp, u64]),
    "gvx_buffer_download": (C.c_int, [u64, u64, vp, u64]),
    "gvx_buffer_device_ptr": (C.c_int, [u64, P(vp), P(u64)]),
    "gvx_map_create": (C.c_int, [P(u64), C.c_int, u32, u32, u32, f32]),
    "gvx_map_create_slab": (C.c_int, [P(u64), C.c_int, u32, u32, u32, f32, u32, u32]),
    "gvx_map_wrap": (C.c_int, [P(u64), C.c_int, vp, u32, u32, u32, f32, u32, u32]),
    "gvx_map_destroy": (C.c_int, [u64]),
    "gvx_map_clear": (C.c_int, [u64]),
    "gvx_map_clear_meaning": (C.c_int, [u64, C.c_int]),
    "gvx_map_clear_bits": (C.c_int, [u64, vp]),
    "gvx_map_device_ptr": (C.c_int, [u64, P(vp), P(u64), P(u64)]),
    "gvx_map_device_ptr_const": (C.c_int, [u64, P(vp), P(u64), P(u64)]),
    "gvx_summary_fallback_count": (u64, []),
    "gvx_map_info": (C.c_int, [u64, P(C.c_int), P(u32), P(f32), P(u32), P(u32), P(u64)]),
    "gvx_map_download": (C.c_int, [u64, vp]),
    "gvx_map_upload": (C.c_int, [u64, vp]),
    "gvx_map_refresh_summary": (C.c_int, [u64]),
    "gvx_map_summary_download": (C.c_int, [u64, vp, u64, P(C.c_int)]),
    "gvx_map_coarse_download": (C.c_int, [u64, vp, u64, P(C.c_int)]),
    "gvx_set_option": (C.c_int, [C.c_int, C.c_int]),
    "gvx_insert_points": (C.c_int, [u64, vp, u64, C.c_int, C.c_int, P(C.c_int)]),
    "gvx_insert_meta": (C.c_int, [u64, vp, C.c_int, vp, u32, vp, u32, vp, P(C.c_int)]),
    "gvx_transform_points": (C.c_int, [vp, vp, u64, C.c_int, vp]),
    "gvx_update_logodds": (C.c_int, [u64, vp, u64, C.c_int, vp, C.c_int]),
    "gvx_insert_sensor_data": (C.c_int, [u64, vp, u64, C.c_int, vp, C.c_int, u64, C.c_int, C.c_int, C.c_int]),
    "gvx_update_logodds_multi": (C.c_int, [u64, vp, C.c_int, vp, u32, vp, C.c_int]),
    "gvx_collide_map_map": (C.c_int, [u64, u64, f32, vp, C.c_int, P(u64)]),
    "gvx_collide_map_map_thresholds": (C.c_int, [u64, u64, C.c_int, C.c_int, vp, C.c_int, P(u64)]),
    "gvx_collide_map_map_async": (C.c_int, [u64, u64, f32, vp, C.c_int, vp]),
    "gvx_collide_map_map_any": (C.c_int, [u64, u64, f32, P(C.c_int)]),
    "gvx_collide_map_map_types": (C.c_int, [u64, u64, f32, C.c_int, vp, P(u64)]),
    "gvx_collide_map_map_bitcheck": (C.c_int, [u64, u64, u32, u32, C.c_int, vp, P(u64)]),
    "gvx_map_shift_sv_ids": (C.c_int, [u64, u32]),
    "gvx_shard_info": (C.c_int, [P(C.c_int), P(C.c_int), P(C.c_int)]),
    "gvx_shard_allreduce_u64": (C.c_int, [vp, C.c_int, C.c_int]),
    "gvx_shard_allgather": (C.c_int, [vp, u32, vp]),
    "gvx_map_ipc_handle": (C.c_int, [u64, vp]),
    "gvx_map_connect_peers": (C.c_int, [u64, C.c_int, C.c_int, vp, vp]),
    "gvx_map_disconnect_peers": (C.c_int, [u64]),
    "gvx_halo_plan": (C.c_int, [u32, u32, u32, vp, C.c_int, C.c_int, vp, vp]),
    "gvx_shard_connect_host": (C.c_int, [C.c_int, C.c_int, C.c_char_p]),
    "gvx_shard_disconnect": (C.c_int, []),
    "gvx_peer_mailbox_create": (C.c_int, [vp]),
    "gvx_peer_mailbox_connect": (C.c_int, [C.c_int, C.c_int, vp]),
    "gvx_peer_mailbox_disconnect": (C.c_int, []),
    "gvx_peer_dispatch_create": (C.c_int, [C.c_int, u64, vp]),
    "gvx_peer_dispatch_connect": (C.c_int, [C.c_int, C.c_int, vp]),
    "gvx_peer_dispatch_disconnect": (C.c_int, []),
    "gvx_update_logodds_dispatch": (C.c_int, [u64, vp, u64, C.c_int, vp, C.c_int, vp]),
    "gvx_collide_map_map_allreduce": (C.c_int, [u64, u64, f32, C.c_int, P(u64), vp]),
    "gvx_collide_list_map_per_meaning_allreduce": (C.c_int, [u64, u64, f32, C.c_int, vp]),
    "gvx_collide_list_map_allreduce": (C.c_int, [u64, u64, f32, C.c_int, P(u64), vp]),
    "gvx_list_create": (C.c_int, [P(u64), u32, u32, u32, f32]),
    "gvx_list_destroy": (C.c_int, [u64]),
    "gvx_list_clear": (C.c_int, [u64]),
    "gvx_list_size": (C.c_int, [u64, P(u64)]),
    "gvx_list_insert_points": (C.c_int, [u64, vp, u64, C.c_int, C.c_int]),
    "gvx_list_insert_meta": (C.c_int, [u64, vp, C.c_int, vp, u32, vp, u32]),
    "gvx_list_download": (C.c_int, [u64, vp, vp, vp]),
    "gvx_list_upload": (C.c_int, [u64, vp, vp, vp, u64]),
    "gvx_list_device_ptrs": (C.c_int, [u64, P(vp), P(vp), P(vp), P(u64)]),
    "gvx_collide_list_map": (C.c_int, [u64, u64, f32, vp, C.c_int, P(u64)]),
    "gvx_collide_list_map_async": (C.c_int, [u64, u64, f32, vp, C.c_int, vp]),
    "gvx_collide_list_map_per_meaning_async": (C.c_int, [u64, u64, f32, vp, C.c_int, vp]),
    "gvx_collide_list_map_types": (C.c_int, [u64, u64, f32, vp, C.c_int, vp, P(u64)]),
    "gvx_collide_list_map_mask": (C.c_int, [u64, u64, f32, vp, C.c_int, vp, P(u64)]),
    "gvx_collide_list_map_per_meaning": (C.c_int, [u64, u64, f32, vp, C.c_int, vp, P(u64)]),
    "gvx_collide_list_list": (C.c_int, [u64, u64, P(u64)]),
    "gvx_collide_list_list_types": (C.c_int, [u64, u64, vp, P(u64)]),
    "gvx_collide_list_list_bitcheck": (C.c_int, [u64, u64, u32, P(u64)]),
    "gvx_collide_list_list_per_meaning": (C.c_int, [u64, u64, vp, P(u64)]),
    "gvx_list_subtract": (C.c_int, [u64, u64]),
    "gvx_list_equals": (C.c_int, [u64, u64, P(C.c_int)]),
    "gvx_list_shift_sv_ids": (C.c_int, [u64, u32]),
    "gvx_map_write_file": (C.c_int, [u64, C.c_char_p]),
    "gvx_map_read_file": (C.c_int, [u64, C.c_char_p]),
    "gvx_list_write_file": (C.c_int, [u64, C.c_char_p]),
    "gvx_list_read_file": (C.c_int, [u64, C.c_char_p]),
    "gvx_list_merge": (C.c_int, [u64, u64, vp, C.c_int]),
    "gvx_pointcloud_file_load": (C.c_int, [C.c_char_p, C.c_int, C.c_int, vp, f32, P(vp), P(u64)]),
    "gvx_pointcloud_file_free": (None, [vp]),
    "gvx_insert_pointcloud_file": (C.c_int, [u64, C.c_char_p, C.c_int, C.c_int, C.c_int, vp, f32, P(C.c_int)]),
    "gvx_list_insert_pointcloud_file": (C.c_int, [u64, C.c_char_p, C.c_int, C.c_int, C.c_int, vp, f32]),
    "gvx_batch_robot_collide": (C.c_int, [u64, vp, C.c_int, vp, u32, vp, vp, u32, vp, f32, vp]),
    "gvx_batch_robot_collide_dh": (C.c_int, [u64, vp, C.c_int, vp, u32, vp, vp, vp, u32, vp, vp, u32, vp, f32, vp]),
    "gvx_dh_chain_transforms": (C.c_int, [vp, vp, u32, vp, u32, vp, vp]),
    "gvx_batch_robot_collide_noalloc": (C.c_int, [u64, vp, vp, u32, vp, vp, u32, vp, f32, vp]),
    "gvx_noalloc_scratch_bytes": (u64, []),
    "gvx_batch_set_scratch_budget": (C.c_int, [u64]),
    "gvx_lane_create": (C.c_int, [vp]),
    "gvx_lane_destroy": (C.c_int, [u64]),
    "gvx_lane_count": (C.c_int, [vp]),
    "gvx_lane_collide_map_map": (C.c_int, [u64, u64, u64, f32, vp, vp]),
    "gvx_lane_collide_map_map_any": (C.c_int, [u64, u64, u64, f32, vp]),
    "gvx_lane_collide_list_map": (C.c_int, [u64, u64, u64, f32, vp, vp]),
}


class GvxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("gvx error %d: %s" % (code, msg))
        self.code = code


_lib = None


def load():
    """dlopen libgvx.so and declare the signatures. No compute call is made."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("libgvx.so is not built (run `python -m gpu_voxels_b200.build`); "
                              "there is no CPU fallback for this path")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(rc):
    if rc != GVX_OK:
        raise GvxError(rc, load().gvx_last_error().decode("utf-8", "replace"))


_initialized = False
_shared_stream = None  # cudaStream_t handed to gvx_set_stream through set_stream(), else None (the library's own stream)


def set_stream(cuda_stream):
    """gvx_set_stream, remembered so that order_after_torch() can skip the event when torch works on the same stream"""
    global _shared_stream
    check(init().gvx_set_stream(C.c_void_p(cuda_stream) if cuda_stream else None))
    _shared_stream = int(cuda_stream) if cuda_stream else None


def order_after_torch(tensor):
    """A CUDA tensor is about to be handed over as GVX_MEM_DEVICE: libgvx enqueues on its own non-blocking stream, which
    has no ordering against torch's streams, so the kernels that read the tensor are made to wait for everything torch
    has queued on its current stream (one event record + one cudaStreamWaitEvent; nothing blocks on the host)."""
    import torch
    cur = torch.cuda.current_stream(tensor.device)
    if _shared_stream is not None and cur.cuda_stream == _shared_stream:
        return
    ev = torch.cuda.Event()
    ev.record(cur)
    check(load().gvx_wait_event(C.c_void_p(ev.cuda_event)))


def init(device=-1):
    global _initialized
    L = load()
    if not _initialized:
        check(L.gvx_init(int(device)))
        _initialized = True
    return L
