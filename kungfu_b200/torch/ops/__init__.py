from .adapt import (calc_stats, check_interference, log_stats, print_strategy_stats,  # noqa: F401
                    resize, resize_cluster, resize_cluster_from_url, set_tree, step_based_schedule)
from .collective import (all_gather, all_reduce_fn, barrier, broadcast,  # noqa: F401
                         broadcast_parameters, consensus, inplace_all_reduce_async_op,
                         inplace_all_reduce_op, inplace_broadcast_async_op,
                         monitored_all_reduce, request_variable, save_variable,
                         wait_all_handles, wait_handle)
from .local import request_model, request_variable_with_template, save_model, save_variables  # noqa: F401
from .monitor import egress_rates, global_noise_scale  # noqa: F401
from .p2p import (AsyncModelAveraging, AsyncRequestModel, ModelAveraging, PeerSelector, RequestModel,  # noqa: F401
                  async_model_averaging, async_request_model, model_averaging)
from .state import counter, exponential_moving_average  # noqa: F401
from .topology import (cluster_size, get_neighbour_mask, get_peer_latencies, get_random_peer,  # noqa: F401
                       global_minimum_spanning_tree, minimum_spanning_tree, mst_to_tree,
                       peer_info, rank, round_robin)


def group_all_reduce(tensors, op='sum'):
    """All-reduce a list of tensors (entries may be None, which stay None)."""
    return [None if t is None else all_reduce_fn(t, op, name='group::%d' % i) for i, t in enumerate(tensors)]


def fuse(tensors):
    """Concatenate flattened tensors (reference: ops/__init__.py:29-31)."""
    import torch
    return torch.cat([t.reshape(-1) for t in tensors])


def defuse(flat, shapes):
    """Inverse of fuse: slice ``flat`` back into tensors of the given ``shapes``."""
    import torch
    out, off = [], 0
    for s in shapes:
        n = int(torch.Size(s).numel())
        out.append(flat[off:off + n].view(s))
        off += n
    return out


# ---- names of the reference's TensorFlow op wrappers (srcs/python/kungfu/tensorflow/ops/collective.py:8-133) ----
def all_reduce(t, op='sum', name=None):
    """``kungfu.tensorflow.ops.all_reduce``: returns the reduced tensor."""
    return all_reduce_fn(t, op, name=name)


def group_nccl_all_reduce(tensors, op='sum'):
    """The reference's NCCL group all-reduce; CUDA tensors here take the peer-memory / NVLS kernels."""
    return group_all_reduce(tensors, op)


def group_hierarchical_nccl_all_reduce(tensors, op='sum'):
    """Local reduce -> cross-host all-reduce -> local broadcast in the reference; the device communicator
    does exactly that by itself when a job spans hosts (kungfu_b200.ops.collective.hier_all_reduce)."""
    return group_all_reduce(tensors, op)
