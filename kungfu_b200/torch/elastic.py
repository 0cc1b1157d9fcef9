"""Elastic training helpers (live state transfer, no checkpoint round trip).

Reference: KungFuElasticTrainHook
(/root/reference/srcs/python/kungfu/tensorflow/hooks/elastic.py:14-87), ElasticHook +
ResizeProfiler (/root/reference/srcs/python/kungfu/tensorflow/experimental/hook/elastic.py:11-116).
After a membership change every member all-reduces (max) the progress counters
and receives the model from rank 0; detached workers stop.
"""
import time

import torch

from .. import python as kfpy
from ..utils import step_based_schedule
from .ops import collective as tops


class ResizeProfiler(object):
    """Durations of resizes as a survivor sees them, in two parts: waiting for the
    new membership (consensus, runners spawning joiners, joiners starting: nothing
    a survivor can speed up) and the state hand-over (re-mapping peer memory,
    re-creating arenas, broadcasting model and optimizer state)."""

    def __init__(self):
        self.records = []
        self.phases = []  # (old, new, membership_s, handover_s)
        self._t0 = None
        self._t1 = None

    def begin(self):
        self._t0 = time.time()
        self._t1 = None

    def membership_done(self):
        self._t1 = time.time()

    def end(self, old, new):
        now = time.time()
        self.records.append((old, new, now - self._t0))
        t1 = self._t1 if self._t1 is not None else self._t0
        self.phases.append((old, new, t1 - self._t0, now - t1))

    def report(self):
        return ['resize %d -> %d took %.3fs (membership %.3fs, state hand-over %.3fs)' % (r + p[2:])
                for r, p in zip(self.records, self.phases)]


def warm_up(model, make_batch, loss_fn, steps=2):
    """Run a few LOCAL forward/backward passes (no optimizer, no communication)
    and drop the gradients: cuDNN's algorithm search, the allocator's first
    growth and lazy module loading then happen before this worker talks to the
    cluster. A worker that joins a running job should call this BEFORE its first
    kungfu call (the control plane starts on first use): the survivors wait for
    a joiner at the membership barrier either way, but the first steps of the new
    cluster then run at full speed instead of waiting for the joiner's autotune."""
    for _ in range(steps):
        x, y = make_batch()
        loss_fn(model(x), y).backward()
    for p in model.parameters():
        p.grad = None
    if torch.cuda.is_available():
        torch.cuda.synchronize()


class ElasticState(object):
    """Tracks step / trained samples and re-synchronises them with the model."""

    def __init__(self, model, optimizer=None, max_step=None):
        self.model, self.optimizer = model, optimizer
        self.step = 0
        self.trained_samples = 0
        self.last_resize_step = 0
        self.max_step = max_step
        self.profiler = ResizeProfiler()
        self.stopped = False
        self.sync()

    def sync(self):
        # everything a policy may base a collective decision on must be equal on
        # survivors and joiners: step, samples AND the step of the last resize
        v = torch.tensor([self.step, self.trained_samples, self.last_resize_step], dtype=torch.int64)
        v = tops.all_reduce_fn(v, 'max', name='kungfu::elastic-progress')
        self.step, self.trained_samples, self.last_resize_step = int(v[0]), int(v[1]), int(v[2])
        if self.optimizer is not None and hasattr(self.optimizer, 'rebuild'):
            # move the flat arenas (parameters, gradients) into the new
            # membership's symmetric memory before anything is broadcast into them
            self.optimizer.rebuild()
        tops.broadcast_parameters(self.model.state_dict())
        if self.optimizer is not None and hasattr(self.optimizer, 'broadcast_state'):
            self.optimizer.broadcast_state()
        # nothing refers to the previous membership's peer memory any more: once every member is
        # past its last use (barrier), unmap the peers' blocks and free ours
        from ..parallel import comm as pc
        if torch.cuda.is_available() and torch.cuda.is_initialized():
            torch.cuda.synchronize()
        kfpy.run_barrier()
        pc.release_retired()

    def end_step(self, batch_size=0):
        self.step += 1
        self.trained_samples += batch_size * kfpy.current_cluster_size()

    def resize(self, new_size):
        """Returns True if this worker keeps training."""
        old = kfpy.current_cluster_size()
        if new_size == old:
            return True
        self.profiler.begin()
        if self.optimizer is not None and hasattr(self.optimizer, 'consolidate'):
            # fused optimizers shard their fp32 state over the ranks: make it whole
            # on every (old) member while all of them are still here
            self.optimizer.consolidate()
        self.last_resize_step = self.step
        changed, detached = kfpy.resize_cluster(new_size)
        self.profiler.membership_done()
        if detached:
            self.stopped = True
            return False
        if changed:
            self.sync()
            self.profiler.end(old, new_size)
        return True

    def should_stop(self):
        return self.stopped or (self.max_step is not None and self.step >= self.max_step)


class ElasticTrainHook(object):
    """Follow a "size:steps,size:steps,..." schedule (KungFuElasticTrainHook)."""

    def __init__(self, schedule, max_step, model, optimizer=None):
        self.schedule = schedule
        self.state = ElasticState(model, optimizer, max_step)

    def after_step(self, batch_size=0):
        self.state.end_step(batch_size)
        want = step_based_schedule(self.schedule, self.state.step, default=kfpy.current_cluster_size(), strict=False)
        return self.state.resize(want) and not self.state.should_stop()


# reference names (tensorflow/hooks/elastic.py:14, tensorflow/experimental/hook/elastic.py:50)
KungFuElasticTrainHook = ElasticTrainHook


class ElasticHook(object):
    """``ElasticHook(local_batch_size, epochs, epoch_size, model)``: the reference's elastic session hook for plain
    training loops. After every step it counts the trained samples, asks the config server for the current cluster
    (``resize_cluster_from_url``, when the job was started with ``-config-server`` / ``-builtin-config-port``) and,
    if the membership changed, re-syncs progress counters and state before the next step. ``after_step()`` returns
    False when this worker should stop: detached by a shrink, or ``epochs * epoch_size`` samples trained."""

    def __init__(self, local_batch_size, epochs, epoch_size, model, optimizer=None):
        import os
        self.local_batch_size = local_batch_size
        self.total_samples = int(epochs * epoch_size)
        self.state = ElasticState(model, optimizer)
        self.exit_reason = None
        self._np = kfpy.current_cluster_size()
        self._poll = bool(os.environ.get('KUNGFU_CONFIG_SERVER'))

    def after_step(self):
        self.state.end_step(self.local_batch_size)
        changed = False
        if self._poll and kfpy.cluster_change_pending():
            self.state.profiler.begin()
            if hasattr(self.state.optimizer, 'consolidate'):
                # sharded fp32 optimizer state is made whole only when a resize is really coming
                # (it used to be gathered after every single step)
                self.state.optimizer.consolidate()
            changed, detached = kfpy.resize_cluster_from_url()
            self.state.profiler.membership_done()
            if detached:
                self.exit_reason = 'change cluster'
                return False
        elif kfpy.detached():
            self.exit_reason = 'change cluster'
            return False
        if changed or kfpy.current_cluster_size() != self._np:
            if not self._poll:
                self.state.profiler.begin()
            old, self._np = self._np, kfpy.current_cluster_size()
            self.state.sync()
            self.state.profiler.end(old, self._np)
        if self.state.trained_samples >= self.total_samples:
            self.exit_reason = 'finished'
            return False
        return True
