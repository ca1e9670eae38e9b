"""ctypes binding of libstgcn_b200.so (the C ABI declared in include/stgcn_b200.h).

There is no CPU fallback: if the library is missing it is built once with nvcc;
if that fails, or a call returns an error code, a RuntimeError is raised.
"""
import ctypes as C
import os
import threading

import torch

from . import _build

F32, BF16 = 0, 1
RES_NONE, RES_IDENTITY, RES_CONV = 0, 1, 2
FLAG_FORCE_SIMT = 1


class BlockDesc(C.Structure):
    _fields_ = [('n', C.c_int32), ('c_in', C.c_int32), ('c_out', C.c_int32), ('t', C.c_int32), ('v', C.c_int32),
                ('k', C.c_int32), ('gamma', C.c_int32), ('stride', C.c_int32), ('pad', C.c_int32), ('dtype', C.c_int32),
                ('training', C.c_int32), ('residual', C.c_int32), ('flags', C.c_int32), ('dropout_p', C.c_float),
                ('bn_eps', C.c_float), ('bn_momentum', C.c_float), ('seed', C.c_uint64), ('offset', C.c_uint64)]


_P = C.c_void_p


class BlockParams(C.Structure):
    _fields_ = [(n, _P) for n in ('bn1_weight', 'bn1_bias', 'bn1_running_mean', 'bn1_running_var',
                                  'bn1_num_batches_tracked', 'A', 'W', 'W_bias', 'Wt', 'Wt_bias', 'bn2_weight',
                                  'bn2_bias', 'bn2_running_mean', 'bn2_running_var', 'bn2_num_batches_tracked',
                                  'Wr', 'Wr_bias', 'rng_state', 'train_prep', 'sync_buf', 'sync_allreduce', 'sync_user')] \
        + [('sync_world', C.c_int64)]

    def set_sync(self, sync):
        """Fill the optional SyncBN fields from a parallel.SyncBatchNormLink (None: leave them zero = local statistics)."""
        if sync is not None:
            self.sync_buf, self.sync_allreduce, self.sync_user, self.sync_world = sync.buf.data_ptr(), sync.fn_ptr, None, sync.world
        return self


class BlockGrads(C.Structure):
    _fields_ = [(n, _P) for n in ('bn1_weight', 'bn1_bias', 'A', 'W', 'W_bias', 'Wt', 'Wt_bias', 'bn2_weight',
                                  'bn2_bias', 'Wr', 'Wr_bias')]


_lib = None
_lock = threading.Lock()


def _declare(lib):
    i32, sz, u64 = C.c_int32, C.c_size_t, C.c_uint64
    D = C.POINTER(BlockDesc)
    lib.stgcn_block_t_out.restype = i32; lib.stgcn_block_t_out.argtypes = [D]
    lib.stgcn_block_saved_bytes.restype = sz; lib.stgcn_block_saved_bytes.argtypes = [D]
    lib.stgcn_block_workspace_bytes.restype = sz; lib.stgcn_block_workspace_bytes.argtypes = [D, i32]
    lib.stgcn_block_forward.restype = i32
    lib.stgcn_block_forward.argtypes = [D, _P, C.POINTER(BlockParams), _P, _P, _P, sz, _P]
    lib.stgcn_colstats_bytes.restype = sz; lib.stgcn_colstats_bytes.argtypes = [i32]
    lib.stgcn_block_forward_chain.restype = i32
    lib.stgcn_block_forward_chain.argtypes = [D, _P, C.POINTER(BlockParams), _P, _P, _P, sz, _P, i32, _P, C.POINTER(i32), _P]
    lib.stgcn_block_prepared_bytes.restype = sz; lib.stgcn_block_prepared_bytes.argtypes = [D]
    lib.stgcn_block_prepare.restype = i32
    lib.stgcn_block_prepare.argtypes = [D, C.POINTER(BlockParams), _P, _P]
    lib.stgcn_block_forward_prepared.restype = i32
    lib.stgcn_block_forward_prepared.argtypes = [D, _P, C.POINTER(BlockParams), _P, _P, _P, sz, _P]
    lib.stgcn_train_prep_bytes.restype = sz; lib.stgcn_train_prep_bytes.argtypes = [D]
    lib.stgcn_train_prep_multi.restype = i32
    lib.stgcn_train_prep_multi.argtypes = [i32, D, C.POINTER(BlockParams), C.POINTER(_P), _P]
    lib.stgcn_block_backward.restype = i32
    lib.stgcn_block_backward.argtypes = [D, _P, _P, C.POINTER(BlockParams), _P, _P, _P, C.POINTER(BlockGrads), _P, sz, _P]
    lib.stgcn_spatial_workspace_bytes.restype = sz; lib.stgcn_spatial_workspace_bytes.argtypes = [D, i32]
    lib.stgcn_spatial_forward.restype = i32
    lib.stgcn_spatial_forward.argtypes = [D, _P, _P, _P, _P, _P, _P, sz, _P]
    lib.stgcn_static_scratch_bytes.restype = sz; lib.stgcn_static_scratch_bytes.argtypes = []
    lib.stgcn_spatial_forward_static.restype = i32
    lib.stgcn_spatial_forward_static.argtypes = [D, _P, _P, _P, _P, _P, _P]
    lib.stgcn_spatial_backward.restype = i32
    lib.stgcn_spatial_backward.argtypes = [D, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, sz, _P]
    lib.stgcn_head_forward.restype = i32
    lib.stgcn_head_forward.argtypes = [i32] * 6 + [_P] * 5 + [_P]
    lib.stgcn_head_backward.restype = i32
    lib.stgcn_head_backward.argtypes = [i32] * 6 + [_P] * 6 + [_P]
    lib.stgcn_cross_entropy_forward.restype = i32
    lib.stgcn_cross_entropy_forward.argtypes = [i32, i32, _P, _P, C.c_int64, _P, _P, _P]
    lib.stgcn_cross_entropy_backward.restype = i32
    lib.stgcn_cross_entropy_backward.argtypes = [i32, i32, _P, _P, _P, _P]
    lib.stgcn_adam_step.restype = i32
    lib.stgcn_adam_step.argtypes = [i32, _P, _P, _P, _P, _P] + [C.c_double] * 5 + [C.c_int64, _P, _P]
    lib.stgcn_collate_pad_augment.restype = i32
    lib.stgcn_collate_pad_augment.argtypes = [i32] * 5 + [_P] * 5 + [_P]
    lib.stgcn_cast.restype = i32; lib.stgcn_cast.argtypes = [i32, i32, _P, _P, sz, _P]
    lib.stgcn_adjacency_normalize.restype = i32
    lib.stgcn_adjacency_normalize.argtypes = [i32, i32, _P, C.c_float, C.c_float, i32, _P, _P]
    lib.stgcn_dropout_mask.restype = i32; lib.stgcn_dropout_mask.argtypes = [u64, u64, C.c_float, sz, _P, _P]
    lib.stgcn_umma_conv_gemm.restype = i32
    lib.stgcn_umma_conv_gemm.argtypes = [i32] * 9 + [_P] * 4 + [_P]
    lib.stgcn_umma_probe.restype = i32
    lib.stgcn_umma_probe.argtypes = [i32] * 4 + [_P] * 3 + [_P]
    lib.stgcn_timing_enable.restype = None; lib.stgcn_timing_enable.argtypes = [i32]
    lib.stgcn_timing_reset.restype = None; lib.stgcn_timing_reset.argtypes = []
    lib.stgcn_timing_classes.restype = i32; lib.stgcn_timing_classes.argtypes = []
    lib.stgcn_timing_class_name.restype = C.c_char_p; lib.stgcn_timing_class_name.argtypes = [i32]
    lib.stgcn_timing_read.restype = i32
    lib.stgcn_timing_read.argtypes = [i32, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
    lib.stgcn_abi_sizeof.restype = sz; lib.stgcn_abi_sizeof.argtypes = [i32]
    for which, struct in enumerate((BlockDesc, BlockParams, BlockGrads)):
        if lib.stgcn_abi_sizeof(which) != C.sizeof(struct):
            raise RuntimeError('stgcn_b200: %s is %d bytes here but %d in libstgcn_b200.so (stale library or binding)'
                               % (struct.__name__, C.sizeof(struct), lib.stgcn_abi_sizeof(which)))
    lib.stgcn_launch_count.restype = u64; lib.stgcn_launch_count.argtypes = []
    lib.stgcn_last_error.restype = C.c_char_p; lib.stgcn_last_error.argtypes = []
    lib.stgcn_version.restype = C.c_char_p; lib.stgcn_version.argtypes = []


def lib():
    """The loaded shared library (built on first use when absent or stale sources are newer)."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                path = _build.LIB
                # built on first use when absent; rebuilt when the sources changed since it was built and a compiler is at hand
                # (on a box without nvcc the shipped library is used as it is)
                if not os.path.exists(path) or (not _build.is_fresh() and _build.have_nvcc()):
                    path = _build.build()
                handle = C.CDLL(path)
                _declare(handle)
                _lib = handle
    return _lib


def check(rc, what):
    if rc != 0:
        raise RuntimeError('stgcn_b200.%s failed (%d): %s' % (what, rc, lib().stgcn_last_error().decode()))


def ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr(device=None):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def dtype_code(dt):
    if dt == torch.float32:
        return F32
    if dt == torch.bfloat16:
        return BF16
    raise RuntimeError('stgcn_b200: unsupported compute dtype %s (float32 or bfloat16)' % dt)


def require_cuda(t, name):
    if not t.is_cuda:
        raise RuntimeError('stgcn_b200: %s must be a CUDA tensor (there is no CPU path)' % name)


# one scratch arena per device, grown on demand; stream-ordered reuse across calls
_ws = {}
_ws_retired = {}     # device -> arenas that were outgrown: kept alive because captured CUDA graphs may still write into them


def workspace(device, nbytes):
    key = (device.index if device.index is not None else torch.cuda.current_device())
    buf = _ws.get(key)
    if buf is None or buf.numel() < nbytes:
        # grow with head-room so steady state never reallocates (CUDA-graph friendly).  An outgrown arena is retired, not
        # freed: a CapturedTrainStep / CapturedForward recorded its address, and replaying that graph after the caching
        # allocator had handed the memory to someone else would corrupt live tensors silently.
        if buf is not None:
            _ws_retired.setdefault(key, []).append(buf)
        buf = torch.empty(int(nbytes * 1.25) + (1 << 20), dtype=torch.uint8, device=device)
        _ws[key] = buf
    return buf


def current_workspace(device):
    """The arena calls on `device` use right now (None before the first call); graph owners keep a reference to it."""
    return _ws.get(device.index if device.index is not None else torch.cuda.current_device())


def release_retired_workspaces():
    """Free the outgrown arenas.  Only safe once every CUDA graph captured before the last growth has been destroyed."""
    _ws_retired.clear()


def launch_count():
    return int(lib().stgcn_launch_count())


def timing_enable(on=True):
    lib().stgcn_timing_enable(int(on))


def timing_reset():
    lib().stgcn_timing_reset()


def timing_read():
    """{class name: (ms, work, launches)} accumulated since the last reset (synchronises the recorded events)."""
    out = {}
    l = lib()
    for i in range(l.stgcn_timing_classes()):
        ms, work, n = C.c_double(), C.c_double(), C.c_uint64()
        l.stgcn_timing_read(i, C.byref(ms), C.byref(work), C.byref(n))
        if n.value:
            out[l.stgcn_timing_class_name(i).decode()] = (ms.value, work.value, int(n.value))
    return out
