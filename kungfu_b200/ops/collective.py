"""Collectives on symmetric buffers (thin launch wrappers around the kernels).

Every function here launches exactly one kernel of csrc/kernels/collectives.cu
on the current stream. Element counts/offsets are in elements of `dtype`.
Reference semantics: /root/reference/srcs/go/kungfu/session/session.go:98-207,
/root/reference/srcs/go/kungfu/session/allreduce.go:11-40.
"""
import ctypes as C

import torch

from . import _cuda as K

_ESIZE = {K.F32: 4, K.BF16: 2, K.F16: 2, K.I32: 4, K.F64: 8, K.I64: 8, K.U8: 1}

# size (bytes) under which the latency-bound one-shot variant wins; refined
# online by kungfu_b200.parallel.chooser from measured bus bandwidth.
ONESHOT_MAX_BYTES = 256 << 10


# large-message variant installed by parallel.planner.VariantChooser (None = built-in preference)
PREFERRED_LARGE = None

# Fault injection for the interference vote (tests, examples/adapt_strategy.py): the variant with
# this code runs on a single CTA, i.e. at a fraction of its bandwidth, until it is set back to None.
# Set it on EVERY rank at the same point: a collective kernel needs the same grid everywhere (CTA b of
# one rank meets CTA b of the others).
SLOW_VARIANT = None


def variant_blocks(comm, algo, channel=0, blocks=None):
    if SLOW_VARIANT is not None and algo == SLOW_VARIANT:
        return 1
    return comm.channel_blocks(channel, blocks)


def choose_algo(comm, nbytes, buf=None, op=K.OP_SUM, dtype=K.F32):
    if comm.np == 1:
        return K.ALGO_ONESHOT
    nvls_ok = (buf is not None and buf.mc_ptr and op == K.OP_SUM and dtype in (K.F32, K.BF16, K.F16))
    if nbytes > ONESHOT_MAX_BYTES and PREFERRED_LARGE is not None:
        if PREFERRED_LARGE != K.ALGO_NVLS or nvls_ok:
            return PREFERRED_LARGE
    if nvls_ok and nbytes > ONESHOT_MAX_BYTES:
        return K.ALGO_NVLS
    return K.ALGO_ONESHOT if nbytes <= ONESHOT_MAX_BYTES else K.ALGO_TWOSHOT


def all_reduce(comm, buf, count, dtype, op='sum', scale=1.0, algo=None, out=None,
               byte_offset=0, out_byte_offset=None, sumsq_out=None, blocks=None,
               threads=None, stream=None, channel=0):
    """out[r][:] = scale * op_r(buf[r][:]) for `count` elements at `byte_offset`.

    TWOSHOT/NVLS may run in place (out is None -> out = buf). ONESHOT always
    needs a distinct `out` (a SymmBuffer; only this rank's copy is written).
    sumsq_out: float32 device tensor with n_local_ranks elements.
    """
    code = dtype if isinstance(dtype, int) else K.dtype_code(dtype)
    opc = op if isinstance(op, int) else K.OPS[op]
    nbytes = count * _ESIZE[code]
    if algo is None:
        algo = choose_algo(comm, nbytes, buf, opc, code)
        if algo == K.ALGO_ONESHOT and out is None:
            algo = K.ALGO_TWOSHOT
    if out is None:
        if algo == K.ALGO_ONESHOT:
            raise ValueError('one-shot all-reduce cannot run in place')
        out = buf
    if out_byte_offset is None:
        out_byte_offset = byte_offset
    in_arr, _k1 = buf.ptr_array(byte_offset)
    out_arr, _k2 = out.ptr_array(out_byte_offset)
    mc_in = (buf.mc_ptr + byte_offset) if buf.mc_ptr else None
    mc_out = (out.mc_ptr + out_byte_offset) if out.mc_ptr else None
    st = getattr(comm, 'stats', None)
    ev0 = st.begin(stream) if st is not None else None
    K.call('kfb_allreduce', comm.channel(channel), algo, in_arr, out_arr, mc_in, mc_out,
           count, code, opc, float(scale),
           sumsq_out.data_ptr() if sumsq_out is not None else None,
           variant_blocks(comm, algo, channel, blocks), threads or comm.default_threads,
           K.stream_ptr(stream))
    if st is not None:
        st.end(ev0, algo, nbytes, stream)
    return out


def broadcast(comm, buf, nbytes, root=0, byte_offset=0, use_multicast=True,
              blocks=None, threads=None, stream=None, channel=0):
    """Every rank's buf[byte_offset : +nbytes] := root's (in place)."""
    out_arr, _k = buf.ptr_array(byte_offset)
    if comm.emulated:
        src = buf.ptrs[root] + byte_offset
    else:
        src = buf.ptrs[comm.rank] + byte_offset
    mc = (buf.mc_ptr + byte_offset) if (buf.mc_ptr and use_multicast) else None
    K.call('kfb_broadcast', comm.channel(channel), root, src, out_arr, mc, nbytes,
           comm.channel_blocks(channel, blocks), threads or comm.default_threads,
           K.stream_ptr(stream))


def all_gather(comm, src, dst, nbytes, src_byte_offset=0, dst_byte_offset=0,
               blocks=None, threads=None, stream=None):
    """dst[r][slot k] := src[k] for all r, k; slot size nbytes (multiple of 16)."""
    out_arr, _k = dst.ptr_array(dst_byte_offset)
    if comm.emulated:
        in_arr, _k2 = src.ptr_array(src_byte_offset)
        inp = C.cast(in_arr, C.c_void_p)
    else:
        inp = src.ptrs[comm.rank] + src_byte_offset
    K.call('kfb_allgather', comm.cstruct, inp, out_arr, nbytes,
           blocks or comm.default_blocks, threads or comm.default_threads,
           K.stream_ptr(stream))


def reduce(comm, buf, out_tensor, count, dtype, op='sum', root=0, byte_offset=0,
           blocks=None, threads=None, stream=None):
    code = dtype if isinstance(dtype, int) else K.dtype_code(dtype)
    in_arr, _k = buf.ptr_array(byte_offset)
    K.call('kfb_reduce', comm.cstruct, root, in_arr, out_tensor.data_ptr(), count,
           code, K.OPS[op] if not isinstance(op, int) else op,
           blocks or comm.default_blocks, threads or comm.default_threads,
           K.stream_ptr(stream))


def gather(comm, buf, out_tensor, nbytes, root=0, byte_offset=0, blocks=None,
           threads=None, stream=None):
    in_arr, _k = buf.ptr_array(byte_offset)
    K.call('kfb_gather', comm.cstruct, root, in_arr, out_tensor.data_ptr(), nbytes,
           blocks or comm.default_blocks, threads or comm.default_threads,
           K.stream_ptr(stream))


def consensus(comm, buf, nbytes, agree_out, byte_offset=0, blocks=None,
              threads=None, stream=None):
    """agree_out (int32 device tensor, n_local_ranks) := 1 iff all ranks' bytes match."""
    in_arr, _k = buf.ptr_array(byte_offset)
    K.call('kfb_consensus', comm.cstruct, in_arr, nbytes, agree_out.data_ptr(),
           blocks or min(comm.default_blocks, 8), threads or comm.default_threads,
           K.stream_ptr(stream))


def edge_reduce(dst_ptr, src_ptr, count, dtype, op='sum', blocks=16, threads=512, stream=None):
    code = dtype if isinstance(dtype, int) else K.dtype_code(dtype)
    K.call('kfb_edge_reduce', dst_ptr, src_ptr, count, code,
           K.OPS[op] if not isinstance(op, int) else op, blocks, threads,
           K.stream_ptr(stream))


def edge_copy(dst_ptr, src_ptr, nbytes, blocks=16, threads=512, stream=None):
    K.call('kfb_edge_copy', dst_ptr, src_ptr, nbytes, blocks, threads, K.stream_ptr(stream))


_HOST_CODE = None
_pinned = {}


def hier_all_reduce(comm, buf, count, dtype, op='sum', scale=1.0, byte_offset=0, stream=None):
    """Hierarchical all-reduce for jobs that span hosts: NVLink kernels inside
    the host, the control plane between host masters, a multicast/push
    broadcast back (reference: KungfuScheduledHierarchicalNcclAllReduce =
    ncclReduce -> CrossAllReduce over TCP -> ncclBroadcast,
    /root/reference/srcs/cpp/src/tensorflow/ops/gpu/collective.cpp:131-153,
    /root/reference/srcs/cpp/src/nccl/controller.cpp:8-40). In place on `buf`."""
    import ctypes as C

    from .. import _host as H
    from .. import python as kfpy
    code = dtype if isinstance(dtype, int) else K.dtype_code(dtype)
    opc = op if isinstance(op, int) else K.OPS[op]
    es = _ESIZE[code]
    nbytes = count * es
    tdt = {K.F32: torch.float32, K.BF16: torch.bfloat16, K.F16: torch.float16, K.I32: torch.int32,
           K.F64: torch.float64, K.I64: torch.int64, K.U8: torch.uint8}[code]
    hdt = {K.F32: H.F32, K.BF16: H.BF16, K.F16: H.F16, K.I32: H.I32, K.F64: H.F64, K.I64: H.I64, K.U8: H.U8}[code]
    if comm.np > 1:
        algo = choose_algo(comm, nbytes, buf, opc, code)
        if algo == K.ALGO_ONESHOT:
            algo = K.ALGO_TWOSHOT
        all_reduce(comm, buf, count, code, opc, scale=1.0, algo=algo, byte_offset=byte_offset, stream=stream)
    if comm.is_local_root:
        dev_view = buf.local[byte_offset:byte_offset + nbytes].view(tdt)
        host = _pinned.get((tdt, count))
        if host is None:
            host = torch.empty(count, dtype=tdt).pin_memory()
            _pinned[(tdt, count)] = host
        s = stream or torch.cuda.current_stream()
        with torch.cuda.stream(s):
            host.copy_(dev_view, non_blocking=True)
        s.synchronize()
        _n = getattr(hier_all_reduce, '_seq', 0) + 1
        hier_all_reduce._seq = _n
        rc = H.lib().kf_cross_all_reduce(kfpy.peer()._h, host.data_ptr(), host.data_ptr(), count, hdt, opc,
                                         b'kungfu::hier')
        if rc != 0:
            raise K.KungFuCudaError('cross-host all-reduce failed: %s' % H.lib().kf_last_error().decode())
        if scale != 1.0 and tdt.is_floating_point:
            host.mul_(scale)
        with torch.cuda.stream(s):
            dev_view.copy_(host, non_blocking=True)
    elif scale != 1.0:
        pass  # the host master's scaled result arrives with the broadcast below
    if comm.np > 1:
        broadcast(comm, buf, (nbytes + 15) // 16 * 16, root=0, byte_offset=byte_offset, stream=stream)
    return buf
