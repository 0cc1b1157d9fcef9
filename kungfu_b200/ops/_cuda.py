"""ctypes binding of libkungfu_cuda.so (the sm_100a kernels + symmetric memory).

The C ABI is declared in include/kungfu_b200.h. The library is built in-tree by
``kungfu_b200._build`` and loaded lazily; on a machine with a GPU a missing
library is an error (there is no silent PyTorch fallback for these ops).
"""
import ctypes as C
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), 'lib', 'libkungfu_cuda.so')

MAX_RANKS = 16
MAX_BLOCKS = 128
PAD_FLAG_WORDS = MAX_BLOCKS * MAX_RANKS
PAD_MISC_WORD = 2 * PAD_FLAG_WORDS
PAD_BYTES = 32768
STATE_EPOCH_WORDS = MAX_BLOCKS
SCRATCH_FLOATS = MAX_BLOCKS + 8
FUSE_SCRATCH_FLOATS = 1024
VAR_WORK_WORDS = 4 * MAX_BLOCKS + 8

F32, BF16, F16, I32, F64, I64, U8 = range(7)
OP_SUM, OP_MIN, OP_MAX, OP_PROD = range(4)
ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_NVLS = range(3)
STEP_PRESIGNALED, STEP_NO_EXIT_BARRIER, STEP_PARTIAL_NORM = 1, 2, 4

DTYPES = {
    torch.float32: F32,
    torch.bfloat16: BF16,
    torch.float16: F16,
    torch.int32: I32,
    torch.float64: F64,
    torch.int64: I64,
    torch.uint8: U8,
}
OPS = {None: OP_SUM, 'sum': OP_SUM, 'min': OP_MIN, 'max': OP_MAX, 'prod': OP_PROD}


class KfbComm(C.Structure):
    _fields_ = [
        ('rank', C.c_int32),
        ('np', C.c_int32),
        ('pads', C.c_void_p * MAX_RANKS),
        ('local_epoch', C.c_void_p),
        ('scratch', C.c_void_p),
        ('mc_pad', C.c_void_p),
    ]


class SgdHyper(C.Structure):
    _fields_ = [
        ('lr', C.c_float),
        ('momentum', C.c_float),
        ('dampening', C.c_float),
        ('weight_decay', C.c_float),
        ('nesterov', C.c_int32),
        ('first_step', C.c_int32),
        # kind 0 = SGD, 1 = AdamW, 2 = Adam (include/kungfu_b200.h)
        ('kind', C.c_int32),
        ('beta1', C.c_float),
        ('beta2', C.c_float),
        ('eps', C.c_float),
        ('bias_corr1', C.c_float),
        ('bias_corr2', C.c_float),
        ('step', C.c_void_p),
        ('exp_avg_sq', C.c_void_p),
    ]


class KungFuCudaError(RuntimeError):
    pass


_lib = None
_lock = threading.Lock()
# number of kernel launches issued through this binding (bench.py reports it)
launch_count = 0

_PTRS = C.POINTER(C.c_void_p)
_I64S = C.POINTER(C.c_int64)

_SIGS = {
    'kfb_last_error': (C.c_char_p, []),
    'kfb_device_info': (C.c_int, [C.c_int] + [C.POINTER(C.c_int)] * 4),
    'kfb_allreduce': (C.c_int, [C.POINTER(KfbComm), C.c_int, _PTRS, _PTRS, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_float, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    'kfb_broadcast': (C.c_int, [C.POINTER(KfbComm), C.c_int, C.c_void_p, _PTRS, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p]),
    'kfb_allgather': (C.c_int, [C.POINTER(KfbComm), C.c_void_p, _PTRS, C.c_int64, C.c_int, C.c_int, C.c_void_p]),
    'kfb_reduce': (C.c_int, [C.POINTER(KfbComm), C.c_int, _PTRS, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    'kfb_gather': (C.c_int, [C.POINTER(KfbComm), C.c_int, _PTRS, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p]),
    'kfb_barrier': (C.c_int, [C.POINTER(KfbComm), C.c_void_p]),
    'kfb_consensus': (C.c_int, [C.POINTER(KfbComm), _PTRS, C.c_int64, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    'kfb_graph_allreduce': (C.c_int, [C.POINTER(KfbComm), _PTRS, C.c_int, C.c_int, C.c_int, _I64S, _I64S, C.c_int, C.c_int,
                                      C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int, C.c_int,
                                      C.c_void_p]),
    'kfb_edge_reduce': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    'kfb_edge_copy': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p]),
    'kfb_fuse': (C.c_int, [_PTRS, _I64S, _I64S, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    'kfb_sum_slots': (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    'kfb_defuse': (C.c_int, [C.c_void_p, C.c_int, _PTRS, _I64S, _I64S, C.c_int, C.c_int, C.c_void_p]),
    'kfb_ssgd_fused_step': (C.c_int, [C.POINTER(KfbComm), _PTRS, C.c_int, _PTRS, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.POINTER(SgdHyper), C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    'kfb_ssgd_fused_step_nvls': (C.c_int, [C.POINTER(KfbComm), C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.POINTER(SgdHyper), C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    'kfb_peer_signal': (C.c_int, [C.POINTER(KfbComm), C.c_int, C.c_void_p]),
    'kfb_set_step_flags': (None, [C.c_int]),
    'kfb_sgd_flat': (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.POINTER(SgdHyper), C.c_int, C.c_int, C.c_void_p]),
    'kfb_sma_fused_step': (C.c_int, [C.POINTER(KfbComm), _PTRS, C.c_int, _PTRS, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float, C.POINTER(SgdHyper), C.c_int, C.c_int, C.c_void_p]),
    'kfb_pair_avg_step_dev': (C.c_int, [C.POINTER(KfbComm), C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.POINTER(SgdHyper), C.c_void_p, C.c_void_p, C.c_ulonglong, C.c_int, C.c_int, C.c_void_p]),
    'kfb_pair_avg_step': (C.c_int, [C.POINTER(KfbComm), C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.POINTER(SgdHyper), C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    'kfb_pair_pull': (C.c_int, [C.POINTER(KfbComm), C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    'kfb_pair_publish': (C.c_int, [C.POINTER(KfbComm), C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_void_p]),
    'kfb_grad_variance': (C.c_int, [C.POINTER(KfbComm), _PTRS, C.c_int, C.c_int64, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    'kfb_sma_fused_step_ex': (C.c_int, [C.POINTER(KfbComm), _PTRS, C.c_int, _PTRS, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float, C.POINTER(SgdHyper), C.c_int, C.c_int, C.c_int, C.c_void_p]),
    'kfb_pair_blend': (C.c_int, [C.POINTER(KfbComm), C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_int, C.c_int, C.c_void_p]),
    'kfb_gns_finalize': (C.c_int, [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_void_p]),
    'kfb_symm_alloc': (C.c_int, [C.c_int, C.c_size_t, C.POINTER(C.c_void_p)]),
    'kfb_symm_export_fd': (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    'kfb_symm_import_fd': (C.c_int, [C.c_int, C.c_int, C.c_size_t, C.POINTER(C.c_void_p)]),
    'kfb_symm_ptr': (C.c_void_p, [C.c_void_p]),
    'kfb_symm_size': (C.c_size_t, [C.c_void_p]),
    'kfb_symm_free': (C.c_int, [C.c_void_p]),
    'kfb_symm_granularity': (C.c_size_t, [C.c_int, C.c_int]),
    'kfb_multicast_supported': (C.c_int, [C.c_int]),
    'kfb_ipc_alloc': (C.c_int, [C.c_int, C.c_size_t, C.POINTER(C.c_void_p), C.c_char_p]),
    'kfb_ipc_open': (C.c_int, [C.c_int, C.c_char_p, C.POINTER(C.c_void_p)]),
    'kfb_ipc_close': (C.c_int, [C.c_void_p]),
    'kfb_ipc_free': (C.c_int, [C.c_void_p]),
    'kfb_mc_create': (C.c_int, [C.c_int, C.c_int, C.c_size_t, C.POINTER(C.c_void_p), C.POINTER(C.c_int)]),
    'kfb_mc_import': (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_size_t, C.POINTER(C.c_void_p)]),
    'kfb_mc_add_device': (C.c_int, [C.c_void_p]),
    'kfb_mc_bind': (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]),
    'kfb_mc_free': (C.c_int, [C.c_void_p]),
    'kfb_close_fd': (C.c_int, [C.c_int]),
    'kfb_maxpool3x3s2_fwd': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    'kfb_maxpool3x3s2_bwd': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    'kfb_gap_fwd': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    'kfb_gap_bwd': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    'kfb_bn_scratch_floats': (C.c_int, [C.c_longlong, C.c_int]),
    'kfb_bn_act_fwd': (C.c_int, [C.c_void_p] * 9 + [C.c_int64, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    'kfb_bn_act_bwd': (C.c_int, [C.c_void_p] * 10 + [C.c_int64, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    'kfb_linear_act': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    'kfb_linear_act_set_mode': (None, [C.c_int]),
    'kfb_ln_partial_floats': (C.c_int, [C.c_longlong, C.c_int]),
    'kfb_colsum_scratch_floats': (C.c_int, [C.c_int]),
    'kfb_colsum': (C.c_int, [C.c_void_p, C.c_longlong, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    'kfb_pack_qkv_grad': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.c_void_p, C.c_void_p]),
    'kfb_add_dropout_ln_fwd': (C.c_int, [C.c_void_p] * 8 + [C.c_longlong, C.c_int, C.c_float, C.c_float, C.c_ulonglong,
                                         C.c_void_p, C.c_uint, C.c_void_p]),
    'kfb_add_dropout_ln_bwd': (C.c_int, [C.c_void_p] * 10 + [C.c_longlong, C.c_int, C.c_float, C.c_ulonglong,
                                         C.c_void_p, C.c_uint, C.c_void_p]),
    'kfb_act_bwd': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]),
}


def available():
    return os.path.exists(LIB_PATH)


def lib():
    """Load (once) and return the ctypes handle; raises if it cannot be loaded."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise KungFuCudaError(
                '%s is missing: run `python -m kungfu_b200._build` (the CUDA '
                'ops have no PyTorch fallback)' % LIB_PATH)
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            try:
                fn = getattr(l, name)
            except AttributeError:
                continue
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(rc):
    if rc != 0:
        raise KungFuCudaError('kungfu_b200 CUDA call failed (%d): %s' %
                              (rc, lib().kfb_last_error().decode()))


def call(name, *args):
    """Invoke an int-returning entry point, raising on error."""
    global launch_count
    launch_count += 1
    check(getattr(lib(), name)(*args))


def stream_ptr(stream=None):
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


def ptr_array(ptrs):
    arr = (C.c_void_p * len(ptrs))(*[int(p) for p in ptrs])
    return C.cast(arr, _PTRS), arr


def i64_array(vals):
    arr = (C.c_int64 * len(vals))(*[int(v) for v in vals])
    return C.cast(arr, _I64S), arr


def i32_array(vals):
    arr = (C.c_int32 * max(1, len(vals)))(*[int(v) for v in vals])
    return C.cast(arr, C.POINTER(C.c_int32)), arr


def dtype_code(dt):
    try:
        return DTYPES[dt]
    except KeyError:
        raise KungFuCudaError('dtype %s is not supported by the CUDA collectives' % dt)


class DevicePtrTensor(object):
    """Expose raw device memory (e.g. a VMM mapping) to torch via __cuda_array_interface__."""

    def __init__(self, ptr, nbytes, owner=None):
        self.ptr = int(ptr)
        self.nbytes = int(nbytes)
        self.owner = owner
        self.__cuda_array_interface__ = {
            'shape': (self.nbytes, ),
            'typestr': '|u1',
            'data': (self.ptr, False),
            'version': 2,
        }


def tensor_from_ptr(ptr, nbytes, device, owner=None):
    """A uint8 torch tensor aliasing [ptr, ptr+nbytes) on `device`."""
    holder = DevicePtrTensor(ptr, nbytes, owner)
    t = torch.as_tensor(holder, device=device)
    t._kfb_owner = holder  # keep the mapping alive as long as the tensor
    return t
