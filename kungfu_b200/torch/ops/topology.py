"""Topology queries. Reference: /root/reference/srcs/python/kungfu/tensorflow/ops/topology.py,
/root/reference/srcs/python/kungfu/tensorflow/ops/__init__.py:49-83,
/root/reference/srcs/cpp/src/tensorflow/ops/cpu/topology.cpp:6-230."""
import ctypes as C

import torch

from ... import _host as H
from ... import python as kfpy
from . import collective as tops


def rank():
    """Rank of this worker (same as kf.current_rank())."""
    return kfpy.current_rank()


def cluster_size():
    """Number of workers (same as kf.current_cluster_size())."""
    return kfpy.current_cluster_size()


def peer_info():
    """(rank, cluster size)."""
    return kfpy.current_rank(), kfpy.current_cluster_size()


def get_peer_latencies():
    """float32[np]: round-trip seconds to every peer (0 for self)."""
    return torch.tensor(kfpy.get_peer_latencies(), dtype=torch.float32)


def minimum_spanning_tree(weights):
    """Prim MST of an n x n weight matrix (symmetrised); returns int32[n-1, 2] edges."""
    w = torch.as_tensor(weights, dtype=torch.float32).contiguous()
    n = w.shape[0]
    edges = (C.c_int32 * max(2 * (n - 1), 1))()
    l = H.lib()
    l.kf_minimum_spanning_tree.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int32)]
    k = l.kf_minimum_spanning_tree(w.data_ptr(), n, edges)
    return torch.tensor(list(edges)[:2 * k], dtype=torch.int32).view(k, 2)


def global_minimum_spanning_tree(self_weights):
    """Every worker passes its row of latencies; everyone gets the same MST edges
    (all-gather -> Prim on every rank: same input, same output)."""
    rows = tops.all_gather(torch.as_tensor(self_weights, dtype=torch.float32).cpu())
    return minimum_spanning_tree(rows)


def mst_to_tree(edges, n):
    """Father array (rooted at 0) for set_tree from MST edges."""
    father = list(range(n))
    for u, v in edges.tolist():
        father[v] = u
    return father


def get_neighbour_mask(edges):
    """bool[np]: which ranks are adjacent to this one in the edge list."""
    n, r = kfpy.current_cluster_size(), kfpy.current_rank()
    mask = torch.zeros(n, dtype=torch.bool)
    for u, v in torch.as_tensor(edges).tolist():
        if u == r:
            mask[v] = True
        if v == r:
            mask[u] = True
    return mask


class RoundRobin(object):
    """Cycle through the indices selected by a mask (peer selection for p2p)."""

    def __init__(self, mask):
        self.choices = [i for i, m in enumerate(torch.as_tensor(mask).tolist()) if m]
        self.pos = 0

    def __call__(self):
        if not self.choices:
            return -1
        c = self.choices[self.pos % len(self.choices)]
        self.pos += 1
        return c


def round_robin(mask):
    """Iterator-like object: each call returns the next index whose ``mask`` entry is true (peer selector of the reference)."""
    return RoundRobin(mask)


def get_random_peer(cluster_size, self_rank, rng=None):
    """Uniform draw from the OTHER ranks: t ~ U[0, np), a self-hit moves to the next rank
    (/root/reference/srcs/python/kungfu/tensorflow/optimizers/async_sgd.py:71-75)."""
    import random
    if cluster_size <= 1:
        return self_rank
    t = (rng or random).randrange(cluster_size)
    return (t + 1) % cluster_size if t == self_rank else t
