"""Pair averaging (AD-PSGD): average with ONE randomly chosen peer per step.

Reference: /root/reference/srcs/python/kungfu/tensorflow/optimizers/async_sgd.py:13-142
 - target ~ U[0, np); if target == rank: (target + 1) % np        (:71-75)
 - first step: publish the model, barrier                           (:106-108)
 - v <- 0.5 (v + v_target); apply local gradients; publish          (:128-142)
With FusedSGD the pull is a one-sided load of the peer's published snapshot out
of its GPU memory (seqlock + double buffer), fused with the blend, the SGD
update and the publication of the new snapshot: one kernel, no barrier, no
NCCL (csrc/kernels/optim.cu:k_pair_avg). Any other optimizer on CUDA
parameters keeps the device path with three pieces: pull + blend kernel, the
wrapped optimizer's own step, publish kernel. Only CPU models and jobs that
span hosts use the host p2p store (request / save), like the reference.
"""
import ctypes as C
import random

import torch

from ... import python as kfpy
from ..ops import collective as tops
from . import core


class _PairAveraging(object):
    def _kf_setup(self):
        self._kf_snap = {}
        self._kf_started = False
        self._kf_rng = random.Random(self._kf_seed)
        self.torn_reads = None
        self._kf_gen = None  # state of the generic device path

    def _pick(self, np_=None, rank=None):
        if np_ is None:
            np_, rank = kfpy.current_cluster_size(), kfpy.current_rank()
        if np_ <= 1:
            return -1
        t = self._kf_rng.randrange(np_)
        return (t + 1) % np_ if t == rank else t

    # -- fused path --------------------------------------------------------------------
    def _fused_init(self):
        from ...ops import _cuda as K
        comm = self.comm
        self.torn_reads = torch.zeros(1, dtype=torch.int32, device=comm.device)
        for g, arenas in self.arenas():
            for a in arenas:
                a.ensure_master(slots=self._n_slots())
                snap = comm.alloc(2 * a.nbytes)
                self._kf_snap[id(a)] = snap
                K.call('kfb_pair_publish', comm.cstruct, a.params_buf.ptrs[max(comm.rank, 0)],
                       snap.ptrs[max(comm.rank, 0)], a.nbytes, a.nbytes,
                       comm.default_blocks, comm.default_threads, K.stream_ptr())
        torch.cuda.current_stream().synchronize()
        kfpy.run_barrier()

    def _fused_step(self):
        from ...ops import _cuda as K
        comm = self.comm
        if comm.cross:
            # The one-sided pull reads the peer's GPU memory; that only reaches the GPUs of this
            # host. A job that spans hosts pairs over the control plane's store, like the reference.
            raise RuntimeError('fused PairAveraging works inside one NVLink domain; this job spans %d hosts: wrap '
                               'a plain torch optimizer instead (host p2p store path)' % kfpy.host_count())
        if len([a for _, ars in self.arenas() for a in ars]) > 1:
            raise RuntimeError('fused PairAveraging needs all parameters in one dtype (one snapshot sequence)')
        if not self._kf_started:
            self._fused_init()
            self._kf_started = True
        me = max(comm.rank, 0)
        device_select = getattr(self, 'device_select', False) and comm.np > 1
        target = -1 if device_select else self._pick(comm.np, comm.rank)
        # one peer pull (NVLink) + a full-model local update (HBM): no CTA-indexed barrier in this kernel, so
        # it takes as many CTAs as the launch check allows instead of the NVLink-sized comm default
        blocks = self._blocks or max(comm.default_blocks, K.MAX_BLOCKS)
        threads = self._threads or comm.default_threads
        for g, arenas in self.arenas():
            for a in arenas:
                h = self._hyper(g, a)
                a.fuse_grads()
                snap = self._kf_snap[id(a)]
                if device_select:
                    # the kernel draws its own peer from (seed, step counter, rank): nothing in this launch
                    # depends on a host-side choice, so the step can be captured in a CUDA graph
                    sa, _keep = K.ptr_array(snap.ptrs)
                    K.call('kfb_pair_avg_step_dev', comm.cstruct, sa, a.nbytes, a.params_buf.ptrs[me], a.code,
                           a.grads_buf.ptrs[me], a.code, a.master.data_ptr(), a.mom.data_ptr(), a.count, 1.0,
                           C.byref(h), self.torn_reads.data_ptr(), self._step_t.data_ptr(),
                           int(self._kf_seed) & 0xFFFFFFFFFFFFFFFF, blocks, threads, K.stream_ptr())
                    continue
                K.call('kfb_pair_avg_step', comm.cstruct, target,
                       snap.ptrs[target] if target >= 0 else snap.ptrs[me], snap.ptrs[me], a.nbytes,
                       a.params_buf.ptrs[me], a.code, a.grads_buf.ptrs[me], a.code,
                       a.master.data_ptr(), a.mom.data_ptr(), a.count, 1.0, C.byref(h),
                       self.torn_reads.data_ptr(), blocks, threads, K.stream_ptr())
        self._after_step()

    # -- any optimizer, CUDA parameters, one NVLink domain: device pull + blend ---------------
    def _device_init(self):
        from ...ops import _cuda as K
        from ...parallel import comm as pc
        from ...parallel.flat import build_arenas
        comm = pc.default_comm()
        arenas = build_arenas(comm, [(n, p) for n, p in self._kf_named], need_grads=False)
        total = sum(a.nbytes for a in arenas)
        offs, o = [], 0
        for a in arenas:
            offs.append(o)
            o += a.nbytes
        gen = {'comm': comm, 'arenas': arenas, 'offs': offs, 'bytes': total,
               'snap': comm.alloc(2 * total),
               'got': torch.zeros(1, dtype=torch.int32, device=comm.device)}
        if len(arenas) > 1:
            gen['pub'] = torch.empty(total, dtype=torch.uint8, device=comm.device)
            gen['pulled'] = torch.empty(total, dtype=torch.uint8, device=comm.device)
        self.torn_reads = torch.zeros(1, dtype=torch.int32, device=comm.device)
        self._kf_gen = gen
        self._device_publish()
        torch.cuda.current_stream().synchronize()
        kfpy.run_barrier()

    def _device_publish(self):
        from ...ops import _cuda as K
        gen = self._kf_gen
        comm, me = gen['comm'], max(gen['comm'].rank, 0)
        if len(gen['arenas']) == 1:
            src = gen['arenas'][0].params_buf.ptrs[me]
        else:
            for a, off in zip(gen['arenas'], gen['offs']):
                gen['pub'][off:off + a.nbytes].copy_(a.params_buf.local[:a.nbytes])
            src = gen['pub'].data_ptr()
        K.call('kfb_pair_publish', comm.cstruct, src, gen['snap'].ptrs[me], gen['bytes'], gen['bytes'],
               K.MAX_BLOCKS, comm.default_threads, K.stream_ptr())

    def _device_step(self):
        from ...ops import _cuda as K
        if self._kf_gen is None:
            self._device_init()
        gen = self._kf_gen
        comm = gen['comm']
        target = self._pick(comm.np, comm.rank)
        if target >= 0:
            snap = gen['snap']
            if len(gen['arenas']) == 1:
                a = gen['arenas'][0]
                K.call('kfb_pair_blend', comm.cstruct, target, snap.ptrs[target], gen['bytes'],
                       a.params_buf.ptrs[comm.rank], a.code, a.count, self.torn_reads.data_ptr(),
                       K.MAX_BLOCKS, comm.default_threads, K.stream_ptr())
            else:
                K.call('kfb_pair_pull', comm.cstruct, target, snap.ptrs[target], gen['bytes'],
                       gen['pulled'].data_ptr(), gen['bytes'], self.torn_reads.data_ptr(), gen['got'].data_ptr(),
                       K.MAX_BLOCKS, comm.default_threads, K.stream_ptr())
                # (the very first pull can find a peer that has not published yet: got == 0, blend weight 0)
                w = gen['got'].to(torch.float32) * 0.5
                for a, off in zip(gen['arenas'], gen['offs']):
                    other = gen['pulled'][off:off + a.nbytes].view(a.dtype)[:a.count]
                    a.flat_params.add_((other - a.flat_params) * w.to(a.dtype))
        super(self.__class__, self).step()
        self._device_publish()

    # -- host path (CPU parameters, or a job that spans hosts): the p2p store ------------------
    def _fused_model(self):
        return torch.cat([tops.to_host(p.data.detach().reshape(-1).float()) for _, p in self._kf_named])

    def _host_native(self):
        """Dense fp32 parameters in host memory: the runtime publishes them by one gather into the store, lands the
        peer's model in a buffer that is kept between steps and blends in one AVX pass (include/kungfu_host.h:
        kf_save_parts, kf_blend_f32); nothing is concatenated, split or cast in Python."""
        return all(p.dtype == torch.float32 and not p.is_cuda and p.data.is_contiguous() for _, p in self._kf_named)

    def _publish(self, name):
        if self._host_native():
            params = [p.data for _, p in self._kf_named]
            kfpy.peer().save_parts(name, [t.data_ptr() for t in params], [4 * t.numel() for t in params])
        else:
            tops.save_variable(self._fused_model(), name)

    def _generic_step(self):
        name = self._kf_model_name
        if not self._kf_started:
            self._publish(name)
            kfpy.run_barrier()
            self._kf_started = True
        target = self._pick()
        if target >= 0 and self._host_native():
            params = [p.data for _, p in self._kf_named]
            total = sum(t.numel() for t in params)
            land = getattr(self, '_kf_landing', None)
            if land is None or land.numel() != total:
                land = self._kf_landing = torch.empty(total, dtype=torch.float32)
            if kfpy.peer().request(int(target), name, land.data_ptr(), total * 4, None):
                kfpy.peer().blend_f32([t.data_ptr() for t in params], [t.numel() for t in params], land.data_ptr())
        elif target >= 0:
            mine = self._fused_model()
            ok, other = tops.request_variable(target, name, mine)
            if ok:
                off = 0
                for _, p in self._kf_named:
                    n = p.numel()
                    o = other[off:off + n].view(p.shape).to(p.device, p.dtype)
                    p.data.add_(o).mul_(0.5)
                    off += n
        super(self.__class__, self).step()
        self._publish(name)

    def _on_device(self):
        if not all(p.is_cuda for _, p in self._kf_named):
            return False
        if kfpy.current_cluster_size() <= 1 or kfpy.host_count() > 1:
            return False
        return True

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if core.is_fused(self):
            self._fused_step()
        elif self._on_device():
            self._device_step()
        else:
            self._generic_step()
        return loss


def PairAveragingOptimizer(optimizer, named_parameters=None, fuse_requests=True,
                           fused_model_name=None, seed=None):
    """AD-PSGD: every step average the model with ONE randomly chosen peer (``v <- (v + v_peer) / 2``), then apply the
    local gradients; no barrier, no all-reduce. With FusedSGD the pull, blend, update and publication are one kernel
    over peer memory; set ``opt.device_select = True`` to let the kernel draw the peer (CUDA-graph capturable).
    ``seed`` seeds the peer choice (default: a function of the rank)."""
    if named_parameters is None:
        named = [('p%d' % i, p) for i, p in enumerate(q for g in optimizer.param_groups for q in g['params'])]
    else:
        named = [(n, p) for n, p in named_parameters if p.requires_grad]
    if seed is None:
        seed = 1000003 * (kfpy.current_rank() + 1)
    return core.wrap(optimizer, _PairAveraging, _kf_named=named,
                     _kf_model_name=fused_model_name or 'kungfu::fused-model', _kf_seed=seed)
