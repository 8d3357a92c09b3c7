"""ctypes binding of libapexb200.so (the C ABI declared in include/apexb200.h).

There is no CPU fallback: if the library is missing or a call fails this module
raises.  Tensors are passed as raw device pointers (``tensor.data_ptr()``) and
the current torch CUDA stream as a ``void*``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("APEXB200_LIB", os.path.join(_HERE, "libapexb200.so"))   # override: A/B of two builds

OK, EINVAL, ECUDA, ENOMEM, ERANGE = 0, -1, -2, -3, -4
LOSS_HALF_SQ, LOSS_HUBER = 0, 1
FMT_U8, FMT_F32 = 0, 1
ST_EMPTY_TREE, ST_BAD_INDEX, ST_BAD_TARGET, ST_BAD_ACTION, ST_PEER_TIMEOUT, ST_QUOTA = 1, 2, 4, 8, 16, 32
IPC_HANDLE_BYTES = 64
FRAC_BITS = 24
LEAF_MAX = (1 << 36) - 1
MAX_CAPACITY = 1 << 24
TD_NOALLOC_MAX_B = 262144
HOST_STEP_SLOTS = 4
BUF_FRAMES, BUF_FIDX, BUF_ACTION, BUF_REWARD, BUF_GAMMA, BUF_QT, BUF_QTPN, BUF_LEAVES = range(8)


class ApexError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("libapexb200: %s (code %d)" % (message, code))
        self.code = code


class ReplayConfig(C.Structure):
    _fields_ = [("capacity", C.c_int64), ("lanes", C.c_int32), ("frames_per_record", C.c_int32),
                ("lookback", C.c_int32), ("frames_per_state", C.c_int32), ("frame_bytes", C.c_int32),
                ("num_actions", C.c_int32), ("alpha", C.c_float), ("device", C.c_int32)]


class ParamSpan(C.Structure):
    _fields_ = [("tensor", C.c_uint64), ("flat_offset", C.c_uint64), ("bytes", C.c_uint64)]


_P, _I32, _I64, _F = C.c_void_p, C.c_int32, C.c_int64, C.c_float

# name -> argument types; every function returns int unless listed in _RESTYPE
SIGNATURES = {
    "apexb200_version": [],
    "apexb200_strerror": [C.c_int],
    "apexb200_last_error": [],
    "apexb200_replay_create": [C.POINTER(_P), C.POINTER(ReplayConfig)],
    "apexb200_replay_destroy": [_P],
    "apexb200_replay_size": [_P, C.POINTER(_I64)],
    "apexb200_replay_device_bytes": [_P, C.POINTER(_I64)],
    "apexb200_replay_status": [_P, C.POINTER(_I32), _P],
    "apexb200_replay_insert": [_P, _I32, _I32, _P, _P, _P, _P, _P, _P, _P, _P, C.POINTER(_I64), _P],
    "apexb200_replay_insert_lanes": [_P, _P, _I32, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "apexb200_tree_build": [_P, _P],
    "apexb200_tree_update": [_P, _P, _P, _I32, _F, _P],
    "apexb200_tree_set_leaves": [_P, _P, _P, _I32, _P],
    "apexb200_tree_total": [_P, _P, _P],
    "apexb200_tree_geometry": [_P, C.POINTER(_I32), C.POINTER(_I64)],
    "apexb200_tree_read": [_P, _I32, _I64, _I64, _P, _P],
    "apexb200_tree_load_leaves": [_P, _P, _I64, _I64, _P],
    "apexb200_sample": [_P, _P, _I32, _P, _P, _P, _P, _P],
    "apexb200_sample_stratified": [_P, C.c_uint64, _I32, _P, _P, _P, _P, _P, _P],
    "apexb200_sample_targets": [_P, _P, _P, _I32, _P, _P, _P, _P],
    "apexb200_gather": [_P, _P, _P, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "apexb200_set_gather_variant": [_I32],
    "apexb200_debug_timeline": [_P, _I32],
    "apexb200_debug_td_timeline": [_P],
    "apexb200_sample_gather": [_P, _P, C.c_uint64, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I32, _P],
    "apexb200_set_pdl": [_I32],
    "apexb200_set_td_cluster": [_I32],
    "apexb200_td_workspace_bytes": [_I32, C.POINTER(_I64)],
    "apexb200_td_fwd": [_P, _P, _P, _P, _P, _P, _I32, _I32, _F, _F, _I32, _F, _P, _P, _P, _P, _P, _P, _I64, _P],
    "apexb200_is_weights": [_P, _P, _I32, _I64, _F, _P, _P],
    "apexb200_td_fwd_noalloc": [_P, _P, _P, _P, _P, _P, _I32, _I32, _F, _F, _I32, _F, _P, _P, _P, _P, _P],
    "apexb200_td_bwd": [_P, _P, _P, _I32, _I32, _I32, _F, _P, _F, _P, _P, _P],
    "apexb200_td_fwd_update": [_P, _P, _P, _P, _P, _P, _P, _P, _I32, _I32, _F, _I32, _F, _P, _P, _P, _P, _P, _P, _I64,
                               _P, _F, _P],
    "apexb200_host_step": [_P, _I32, _P, _P, _I32, _I32, _F, _I32, _F, _P, _P, _P, _P, _P, _P, _P, _P],
    "apexb200_host_step_sharded": [_P, _I32, C.c_uint64, _I32, _I32, _P, _I32, _F, _I32, _F, _P, _P, _P, _P, _P, _P, _P, _P],
    "apexb200_nstep": [_P, _P, _I32, _I32, _I32, _F, _P, _P, _P, _P],
    "apexb200_initial_priority": [_P, _P, _P, _P, _P, _I32, _I32, _P, _P],
    "apexb200_ingest_create": [_P, _I32, _P, _I32, _F, _I32, _I32, C.POINTER(_P)],
    "apexb200_ingest_destroy": [_P],
    "apexb200_ingest_push": [_P, _P, _I32, _I32, _P, _P, _P, _P, _P, _P],
    "apexb200_ingest_push_blobs": [_P, _P, _P, _I32, _I32, _I32, _I32, _I32, _I32, _P],
    "apexb200_ingest_sync_heads": [_P, _P],
    "apexb200_host_register_mapped": [_P, C.c_uint64, C.POINTER(_P)],
    "apexb200_shard_targets": [_P, _I32, _I32, _P, _I32, _P, _P, _P, _P],
    "apexb200_peer_export": [_P, _P],
    "apexb200_peer_connect": [_P, _I32, _I32, _P],
    "apexb200_peer_connect_local": [_P, _I32, _I32, _P],
    "apexb200_shard_exchange": [_P, _P, _P],
    "apexb200_shard_push": [_P, _P],
    "apexb200_peer_push_on_update": [_P, _I32],
    "apexb200_peer_wait_stats": [_P, _P, _I32, _P],
    "apexb200_sample_gather_sharded": [_P, C.c_uint64, _I32, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I32, _P],
    "apexb200_params_create": [_I32, _P, _I32, C.c_uint64, C.POINTER(_P)],
    "apexb200_params_destroy": [_P],
    "apexb200_params_flat": [_P, C.POINTER(_P), C.POINTER(C.c_uint64)],
    "apexb200_params_publish": [_P, _P, _P, _I64, _P],
    "apexb200_params_load": [_P, _P, _I32, _P],
    "apexb200_actor_policy": [_P, _P, _P, _P, _P, C.c_uint64, C.c_uint64, _I32, _I32, _P, _P, _P, _P],
    "apexb200_host_register": [_P, C.c_uint64],
    "apexb200_host_unregister": [_P],
    "apexb200_replay_buffer": [_P, _I32, C.POINTER(_P), C.POINTER(_I64)],
    "apexb200_frames_pack_sizes": [_P, _I64, _I32, _P, _P],
    "apexb200_frames_pack": [_P, _I64, _I32, _P, _P, _P],
    "apexb200_frames_unpack": [_P, _I64, _I32, _P, _P, _P],
    "apexb200_replay_get_heads": [_P, C.POINTER(_I64)],
    "apexb200_replay_set_heads": [_P, C.POINTER(_I64)],
}
_RESTYPE = {"apexb200_strerror": C.c_char_p, "apexb200_last_error": C.c_char_p}

_lib = None


def lib():
    """Load the shared library once; raise if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "%s not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or ape-x-dqn_b200/csrc/build.sh (there is no CPU fallback)" % LIB_PATH)
        handle = C.CDLL(LIB_PATH)
        for name, argtypes in SIGNATURES.items():
            fn = getattr(handle, name)            # AttributeError if the .so is stale
            fn.argtypes = argtypes
            fn.restype = _RESTYPE.get(name, C.c_int)
        _lib = handle
    return _lib


def check(code):
    if code != OK:
        raise ApexError(code, lib().apexb200_last_error().decode() or lib().apexb200_strerror(code).decode())


def call(name, *args):
    check(getattr(lib(), name)(*args))


def ptr(t):
    """Device (or host) pointer of a tensor / None -> NULL."""
    if t is None:
        return None
    return C.c_void_p(t.data_ptr())


def stream_ptr(stream=None):
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


def host_register(address, nbytes):
    """Page-lock [address, address + nbytes) for this process so that copies to and from it are asynchronous."""
    call("apexb200_host_register", C.c_void_p(address), int(nbytes))


def host_register_mapped(address, nbytes):
    """Page-lock [address, address + nbytes) and map it for kernels; returns the device-visible address."""
    dp = C.c_void_p()
    call("apexb200_host_register_mapped", C.c_void_p(address), int(nbytes), C.byref(dp))
    return int(dp.value)


def host_unregister(address):
    call("apexb200_host_unregister", C.c_void_p(address))


def open_shared_memory(name, create, size=0):
    """multiprocessing.shared_memory.SharedMemory whose lifetime is explicit: the segment is taken off the
    resource tracker, which would otherwise unlink it when *any* process that merely attached to it (or the
    learner that created it) exits — rings and boards outlive the processes that come and go around them.
    The owner unlinks it with .unlink()."""
    from multiprocessing import resource_tracker, shared_memory
    shm = shared_memory.SharedMemory(name=name, create=create, size=size if create else 0)
    try:
        resource_tracker.unregister(shm._name, "shared_memory")
    except Exception:
        pass
    return shm


def unlink_shared_memory(shm):
    """shm.unlink() without the resource-tracker bookkeeping that open_shared_memory removed."""
    import _posixshmem
    try:
        _posixshmem.shm_unlink(shm._name)
    except FileNotFoundError:
        pass
