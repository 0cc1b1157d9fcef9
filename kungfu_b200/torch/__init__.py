"""kungfu_b200.torch — the reference's ``kungfu.torch`` surface
(/root/reference/srcs/python/kungfu/torch/__init__.py:1-14) plus everything the
reference only offered for TensorFlow."""
from ..python import (_get_cuda_index, current_cluster_size, current_local_rank,  # noqa: F401
                      current_local_size, current_rank, detached, propose_new_size,
                      resize_cluster, resize_cluster_from_url, run_barrier, uid)
from . import ops, optimizers  # noqa: F401

get_cuda_index = _get_cuda_index


def nccl_built():
    """Kept for API parity. The GPU path here does not use NCCL at all: CUDA
    tensors are reduced by hand-written kernels over NVLink peer memory."""
    return False


broadcast_parameters = ops.broadcast_parameters
