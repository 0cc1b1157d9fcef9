"""Process-level API: rank / size / barrier / elastic control.

Same surface as the reference's ``kungfu.python``
(/root/reference/srcs/python/kungfu/python/__init__.py:35-107), backed by the
C++ control plane in libkungfu_host.so (csrc/ctrl, csrc/elastic). A process
that was not started by ``kungfu-run`` (or torchrun) runs in single mode:
rank 0 of 1 (/root/reference/srcs/go/kungfu/env/config.go:59-68).
"""
import atexit
import os

from .. import _host

_peer = None


def _p():
    global _peer
    if _peer is None:
        _peer = _host.Peer.from_env()
        atexit.register(_finalize)
    return _peer


def _finalize():
    global _peer
    if _peer is not None:
        _peer.close()
        _peer = None


def init():
    """Start the control plane now (otherwise it starts on first use)."""
    _p()


def finalize():
    """Stop the control plane (also registered with atexit)."""
    _finalize()


def uid():
    """ip:port:init-version packed into an int, unique per worker incarnation."""
    return _p().uid()


def detached():
    """True once this worker is no longer part of the cluster (after a shrink)."""
    return _p().detached()


def current_rank():
    """Rank of this worker in the current cluster, 0 .. size-1 (changes after a resize)."""
    return _p().rank()


def current_local_rank():
    """Rank among the workers of this host (the GPU index by default)."""
    return _p().local_rank()


def current_cluster_size():
    """Number of workers in the current cluster (changes after a resize)."""
    return _p().size()


def current_local_size():
    """Number of workers on this host."""
    return _p().local_size()


def host_count():
    """Number of hosts the current cluster spans."""
    return _p().host_count()


def run_barrier():
    """Block until every worker of the current cluster has called it."""
    _p().barrier()


def propose_new_size(new_size):
    """Ask the config server for a new cluster size (takes effect at the next
    resize_cluster_from_url on every worker)."""
    _p().propose_new_size(int(new_size))


def resize_cluster(new_size):
    """Returns (changed, detached). Surviving workers are NOT restarted: the
    runtime re-connects the control plane and the device layer re-maps peer
    memory (reference: /root/reference/srcs/go/kungfu/peer/peer.go:227-263)."""
    changed, det = _p().resize_cluster(int(new_size))
    if changed:
        _on_membership_change(det)
    return changed, det


def resize_cluster_from_url():
    """Adopt the cluster the config server currently publishes (-config-server / -builtin-config-port).
    Returns (changed, detached) like resize_cluster."""
    changed, det = _p().resize_cluster_from_url()
    if changed:
        _on_membership_change(det)
    return changed, det


def cluster_change_pending():
    """True if the config server publishes a cluster other than the current one, i.e. the next
    resize_cluster_from_url() will change the membership. Collective (every worker gets the same
    answer); nothing is changed. Lets a training loop prepare for a resize only when one is coming."""
    return _p().peek_cluster_change()


def _on_membership_change(detached=False):
    try:
        from ..torch.ops import collective as tops
        tops._reset_names()
        tops.reset_staging()
        from ..parallel import comm
        comm.reset_default_comm()
        if detached:
            # a worker that left the cluster gives its peer mappings and blocks back right away:
            # the survivors' GPUs hold imports of this memory until both sides have let go
            comm.shutdown()
    except Exception:  # device layer not in use
        pass


def _device_planner():
    """The device data plane's planner, if this process has created a device communicator."""
    try:
        from ..parallel import comm as pc
        if pc._default is None:
            return None
        from ..parallel.planner import planner_for
        return planner_for(pc._default)
    except Exception:
        return None


def check_interference():
    """Majority vote over the workers: has the active strategy dropped well below its reference throughput?
    (reference: session/adaptiveStrategies.go:18-127). Two data planes are voted on: the host runtime's
    strategy graphs (CPU tensors) and, when GPU collectives are in use, the device kernel variants fed by
    the training all-reduces; on a device vote the slow variant is replaced by the next best one. The
    return value keeps the reference's meaning for the caller: True = the workers agreed that the
    active strategy is being interfered with."""
    host = _p().check_interference()
    dp = _device_planner()
    dev = dp.check_interference() if dp is not None else False
    return bool(host) or bool(dev)


def calc_stats():
    """Fold the traffic counters since the last call into the per-strategy throughput statistics
    (host strategies and device kernel variants)."""
    _p().calc_stats()
    dp = _device_planner()
    if dp is not None:
        dp.calc_stats()


def log_stats():
    """Log the per-strategy throughput statistics."""
    _p().log_stats()
    dp = _device_planner()
    if dp is not None:
        for line in dp.comm.stats.lines():
            print('[kungfu] rank %d %s' % (current_rank(), line), flush=True)


def print_strategy_stats():
    """Print the per-strategy throughput statistics of this worker."""
    _p().print_strategy_stats()
    dp = _device_planner()
    if dp is not None:
        for line in dp.comm.stats.lines():
            print(line, flush=True)


def set_tree(forest):
    """Install a broadcast tree given as a father array (forest[i] = parent of
    i, root points to itself) on every rank after digest consensus."""
    _p().set_tree([int(x) for x in forest])


def set_strategy(name):
    """Switch every worker to a named strategy (STAR, RING, CLIQUE, TREE, BINARY_TREE, BINARY_TREE_STAR,
    MULTI_BINARY_TREE_STAR, AUTO) after consensus."""
    _p().set_strategy(name)


def get_peer_latencies():
    """Round-trip ping time to every peer in seconds (0 for self): the input of the minimum-spanning-tree strategy."""
    return _p().peer_latencies()


def egress_rates():
    """Bytes per second sent to every peer: the control plane's sockets over the last monitoring period
    (KUNGFU_CONFIG_ENABLE_MONITORING) plus, when GPU collectives are in use, the NVLink bytes the device
    kernels moved since the previous call (per-variant byte model, parallel/devstats.py)."""
    rates = list(_p().egress_rates())
    dp = _device_planner()
    if dp is not None and dp.comm.np == len(rates) and not dp.comm.cross:
        rates = [a + b for a, b in zip(rates, dp.comm.stats.egress_rates())]
    return rates


def _cuda_index_from_env(kf_visible, visible, local_rank_env, local_rank):
    """Pure function behind get_cuda_index() (testable without touching os.environ)."""
    if kf_visible not in (None, ''):
        first = int(kf_visible.split(',')[0])
        if first < 0:
            return -1
        if visible:
            toks = [t.strip() for t in visible.split(',') if t.strip()]
            try:
                ids = [i for i in (int(t) for t in toks) if i >= 0]
            except ValueError:
                ids = None  # GPU-<uuid> / MIG-... entries: the launcher passes the slot
            if ids is not None and first in ids:
                return ids.index(first)  # the launcher names the device; the runtime counts within the visible list
            if len(toks) == 1:
                return 0  # already narrowed to one device
            return first if first < len(toks) else -1
        return first
    if local_rank_env not in (None, ''):
        return int(local_rank_env)
    return local_rank()


def _get_cuda_index():
    """GPU (index for torch.cuda.set_device) this worker should use. The launcher
    names a device in KUNGFU_CUDA_VISIBLE_DEVICES; it is looked up in
    CUDA_VISIBLE_DEVICES, which the launcher narrows to that one device unless
    -allow-nvlink keeps the whole node visible
    (/root/reference/srcs/cpp/src/python/cuda.cpp:29-47,
    /root/reference/srcs/go/kungfu/job/cuda_visible_device.go:17-58). Without a
    KungFu launcher: LOCAL_RANK, else the local rank."""
    return _cuda_index_from_env(os.environ.get('KUNGFU_CUDA_VISIBLE_DEVICES'), os.environ.get('CUDA_VISIBLE_DEVICES'),
                                os.environ.get('LOCAL_RANK'), current_local_rank)


def _get_other_ranks():
    self_rank = current_rank()
    return [r for r in range(current_cluster_size()) if r != self_rank]


def show_cuda_version():
    """Print the CUDA version PyTorch was built with."""
    import torch
    print('CUDA (torch build): %s' % torch.version.cuda)


def show_nccl_version():
    """Print the NCCL version bundled with PyTorch (NCCL is only the comparison baseline here)."""
    import torch
    try:
        print('NCCL (torch bundled, baseline only): %s' % '.'.join(map(str, torch.cuda.nccl.version())))
    except Exception:
        print('NCCL: unavailable')


def control_exchange():
    """The control-plane exchange object the device layer bootstraps from."""
    return _p().exchange()


def peer():
    """The process-wide kungfu_b200._host.Peer (started on first use)."""
    return _p()
